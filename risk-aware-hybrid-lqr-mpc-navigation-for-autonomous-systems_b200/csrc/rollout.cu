// Closed-loop rollouts on the device (BASELINE config 5 and device-side replays of
// configs 1/2): R independent robots follow a shared reference table for `ticks` plant steps.
//
//   reference loop                                   here
//   run_lqr_simulation     run_simulation.py:58-96   mode B2C_ROLL_LQR    : LQR law every tick
//   run_mpc_simulation     :225-280                  mode B2C_ROLL_MPC    : solve_with_ltv every mpc_rate ticks, hold
//   run_hybrid_simulation  :500-570                  mode B2C_ROLL_HYBRID : risk -> hysteresis -> LQR | MPC
//   RiskMetrics.assess_risk (distance risk only, as the hybrid loop calls it: HC/controllers/risk_metrics.py:84-222)
//   DifferentialDriveRobot.simulate_step             HC/models/differential_drive.py:138-172 (clip, Euler, wrap)
//
// Per tick:  k_roll_decide (every robot: reference row, risk, hysteresis; LQR robots get their control from the
// per-row gain table, MPC robots are appended to a work list)  ->  k_roll_gather (work list -> state, obstacles,
// start row of the QP: the table form of K3 gathers the reference segment itself)  ->  the K3 kernels, unchanged,
// which read the length of the work list FROM DEVICE MEMORY  ->  k_roll_scatter  ->  k_roll_plant (every robot).
// Nothing in the loop waits for the host: a 500-tick rollout is 500 x 7 launches enqueued back to back.
// The LQR gain depends on the table row only (compute_gain(u_ref[0], x_ref[2]), lqr_controller.py:191-215), so
// the DARE runs once per table row per call (K2), not once per robot per tick.
#include <vector>
#include "mpc_args.cuh"
#define B2C_GEN_MAX_OBS 8

namespace {

struct RollArgs {
    int64_t R;
    int mode, t, mpc_rate, min_dwell, rows, M, N;
    double dt, plant_v, plant_w, lqr_v, lqr_w;
    double d_safe_risk, d_trigger, alpha, threshold_low, d_safe_mpc;
    const double* table;        // (rows, 6) [t, px, py, th, v, w]
    const double* ktab;         // [6][rows] LQR gain per table row
    const int32_t* k0;          // [R]
    double* x;                  // [3][R]
    const double* obs;          // [3M][R]
    double* u;                  // [2][R] control applied this tick (held between MPC solves)
    int32_t* prev;              // [R] -1 none, 0 LQR, 1 MPC
    int32_t* since;             // [R] steps since the last switch
    int32_t* counters;          // [5][R] lqr_steps, mpc_steps, switches, collision_states, mpc_fallbacks
    double* stats;              // [3][R] sum |e_pos|, max |e_pos|, last |e_pos|
    int32_t* work;              // [R] robots that solve a QP this tick
    unsigned long long* nwork;
    // optional per-tick logs
    double* log_x; double* log_u; double* log_risk; int8_t* log_ctrl;
};

B2C_HD int roll_row(const RollArgs& g, int64_t r, int ahead) {
    const int k = g.k0[r] + g.t + ahead;
    return k < g.rows - 1 ? k : g.rows - 1;          // get_reference_at_index / get_trajectory_segment clamp
}

// RiskMetrics.compute_distance_risk (:84-132) folded into assess_risk with predicted_states=None (:188-222)
B2C_HD double roll_risk(const RollArgs& g, int64_t r, double px, double py) {
    double max_risk = 0.0;
    for (int m = 0; m < g.M; ++m) {
        const double ox = g.obs[(int64_t)(3 * m) * g.R + r], oy = g.obs[(int64_t)(3 * m + 1) * g.R + r];
        const double rad = g.obs[(int64_t)(3 * m + 2) * g.R + r];
        const double dx = px - ox, dy = py - oy;
        const double edge = sqrt(B2C_MUL(dx, dx) + B2C_MUL(dy, dy)) - rad;
        double risk;
        if (edge <= g.d_safe_risk) risk = 1.0;
        else if (edge >= g.d_trigger) risk = 0.0;
        else risk = 1.0 - (edge - g.d_safe_risk) / (g.d_trigger - g.d_safe_risk);
        max_risk = fmax(max_risk, risk);
    }
    return B2C_MUL(g.alpha, max_risk);               // + beta * 0 (the hybrid loop passes no prediction)
}

// one robot, first half of a tick; returns true when the robot needs a QP solve
B2C_HD bool roll_decide_one(const RollArgs& g, int64_t r) {
    const int64_t R = g.R;
    const double px = g.x[r], py = g.x[R + r], th = g.x[2 * R + r];
    const double* row = g.table + (int64_t)roll_row(g, r, 0) * 6;
    const double rx = row[1], ry = row[2], rth = row[3];
    double risk = 0.0;
    bool use_mpc;
    if (g.mode == B2C_ROLL_LQR) use_mpc = false;
    else if (g.mode == B2C_ROLL_MPC) use_mpc = true;
    else {                                            // run_simulation.py:529-549
        risk = roll_risk(g, r, px, py);
        const bool rec = risk >= g.threshold_low;
        const int prev = g.prev[r];
        int since = g.since[r];
        use_mpc = (since >= g.min_dwell || prev < 0) ? rec : (prev == 1);
        const int cur = use_mpc ? 1 : 0;
        if (prev >= 0 && cur != prev) { g.counters[2 * R + r] += 1; since = 0; } else since += 1;
        g.prev[r] = cur; g.since[r] = since;
    }
    // tracking error statistics of this tick (position part, as run_simulation.py:563-566 / :97-100 report)
    {
        const double ex = px - rx, ey = py - ry;
        const double en = sqrt(B2C_MUL(ex, ex) + B2C_MUL(ey, ey));
        g.stats[r] += en;
        g.stats[R + r] = fmax(g.stats[R + r], en);
        g.stats[2 * R + r] = en;
    }
    if (g.log_risk) g.log_risk[(int64_t)g.t * R + r] = risk;
    if (g.log_ctrl) g.log_ctrl[(int64_t)g.t * R + r] = use_mpc ? 1 : 0;
    if (!use_mpc) {                                   // lqr_controller.py:149-215 with the row's gain
        const int kr = roll_row(g, r, 0);
        const double e0 = px - rx, e1 = py - ry, e2 = b2c_wrap(th - rth);
        const double* K = g.ktab + kr;
        const int64_t S = g.rows;
        const double f0 = -K[0] * e0 + -K[S] * e1 + -K[2 * S] * e2;
        const double f1 = -K[3 * S] * e0 + -K[4 * S] * e1 + -K[5 * S] * e2;
        g.u[r] = b2c_clip(row[4] + f0, -g.lqr_v, g.lqr_v);
        g.u[R + r] = b2c_clip(row[5] + f1, -g.lqr_w, g.lqr_w);
        g.counters[r] += 1;
        return false;
    }
    if (g.mode == B2C_ROLL_MPC && (g.t % g.mpc_rate) != 0) return false;    // hold, :250-256
    g.counters[R + r] += 1;
    return true;
}

// second half: DifferentialDriveRobot.simulate_step (clip, explicit Euler, wrap) + collision census
B2C_HD void roll_plant_one(const RollArgs& g, int64_t r) {
    const int64_t R = g.R;
    const double px = g.x[r], py = g.x[R + r], th = g.x[2 * R + r];
    const double u0 = g.u[r], u1 = g.u[R + r];
    if (g.log_u) { g.log_u[((int64_t)g.t * 2) * R + r] = u0; g.log_u[((int64_t)g.t * 2 + 1) * R + r] = u1; }
    const double v = b2c_clip(u0, -g.plant_v, g.plant_v), w = b2c_clip(u1, -g.plant_w, g.plant_w);
    double sn, cs; sincos(th, &sn, &cs);
    const double nx = px + B2C_MUL(g.dt, B2C_MUL(v, cs));
    const double ny = py + B2C_MUL(g.dt, B2C_MUL(v, sn));
    const double nth = b2c_wrap(th + B2C_MUL(g.dt, w));
    g.x[r] = nx; g.x[R + r] = ny; g.x[2 * R + r] = nth;
    if (g.log_x) {
        double* o = g.log_x + ((int64_t)(g.t + 1) * 3) * R + r;
        o[0] = nx; o[R] = ny; o[2 * R] = nth;
    }
    bool hit = false;                                 // Obstacle.is_collision, mpc_controller.py:43-46
    for (int m = 0; m < g.M; ++m) {
        const double dx = nx - g.obs[(int64_t)(3 * m) * R + r], dy = ny - g.obs[(int64_t)(3 * m + 1) * R + r];
        hit = hit || (sqrt(B2C_MUL(dx, dx) + B2C_MUL(dy, dy)) < g.obs[(int64_t)(3 * m + 2) * R + r] + g.d_safe_mpc);
    }
    if (hit) g.counters[3 * R + r] += 1;
}

// QP inputs of work item j (robot r): x0 = state, first reference row (the table form of K3 takes rows k..k+N of
// the shared table itself, clamped at its end like get_trajectory_segment), the robot's obstacles
B2C_HD void roll_gather_one(const RollArgs& g, int64_t j, int64_t cap, double* x0, int32_t* qk0,
                            double* ob, int32_t* sc, const int32_t* step_count) {
    const int64_t r = g.work[j], R = g.R;
    x0[j] = g.x[r]; x0[cap + j] = g.x[R + r]; x0[2 * cap + j] = g.x[2 * R + r];
    qk0[j] = roll_row(g, r, 0);
    for (int e = 0; e < 3 * g.M; ++e) ob[(int64_t)e * cap + j] = g.obs[(int64_t)e * R + r];
    sc[j] = step_count[r];
}

B2C_HD void roll_scatter_one(const RollArgs& g, int64_t j, int64_t cap, const double* u0, const int32_t* sc,
                             const int32_t* st, int32_t* step_count) {
    const int64_t r = g.work[j], R = g.R;
    g.u[r] = u0[j]; g.u[R + r] = u0[cap + j];
    step_count[r] = sc[j];
    if (st[j] != B2C_OK) g.counters[4 * R + r] += 1;
}

#ifndef B2C_HOST_EMUL
__global__ void __launch_bounds__(256) k_roll_decide(const RollArgs g) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool need = (r < g.R) && roll_decide_one(g, r);
    const unsigned mask = __ballot_sync(0xffffffffu, need);
    if (!mask) return;
    const int lane = threadIdx.x & 31, leader = __ffs(mask) - 1;
    unsigned long long base = 0;
    if (lane == leader) base = atomicAdd(g.nwork, (unsigned long long)__popc(mask));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (need) g.work[base + __popc(mask & ((1u << lane) - 1u))] = (int32_t)r;
}
__global__ void __launch_bounds__(256) k_roll_plant(const RollArgs g) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r < g.R) roll_plant_one(g, r);
}
// the work list's length is read on the device (no host round trip per tick); the grid covers all R robots
__global__ void __launch_bounds__(256)
k_roll_gather(const RollArgs g, int64_t cap, double* x0, int32_t* qk0, double* ob, int32_t* sc, const int32_t* step_count) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j < (int64_t)*g.nwork) roll_gather_one(g, j, cap, x0, qk0, ob, sc, step_count);
}
__global__ void __launch_bounds__(256)
k_roll_scatter(const RollArgs g, int64_t cap, const double* u0, const int32_t* sc, const int32_t* st,
               int32_t* step_count) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j < (int64_t)*g.nwork) roll_scatter_one(g, j, cap, u0, sc, st, step_count);
}
// (the work list is filled with atomics, so its order varies from launch to launch; nothing a robot computes
// depends on its position in the list)
__global__ void __launch_bounds__(256) k_roll_table_cols(int rows, const double* table, double* v, double* th) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < rows) { v[k] = table[(int64_t)k * 6 + 4]; th[k] = table[(int64_t)k * 6 + 3]; }
}
__global__ void __launch_bounds__(256) k_roll_init(const RollArgs g) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= g.R) return;
    g.prev[r] = -1; g.since[r] = 0; g.u[r] = 0.0; g.u[g.R + r] = 0.0;
    for (int c = 0; c < 5; ++c) g.counters[(int64_t)c * g.R + r] = 0;
    for (int c = 0; c < 3; ++c) g.stats[(int64_t)c * g.R + r] = 0.0;
    if (g.log_x) { g.log_x[r] = g.x[r]; g.log_x[g.R + r] = g.x[g.R + r]; g.log_x[2 * g.R + r] = g.x[2 * g.R + r]; }
    bool hit = false;                                 // the initial state is part of the collision census (:288-293)
    for (int m = 0; m < g.M; ++m) {
        const double dx = g.x[r] - g.obs[(int64_t)(3 * m) * g.R + r], dy = g.x[g.R + r] - g.obs[(int64_t)(3 * m + 1) * g.R + r];
        hit = hit || (sqrt(dx * dx + dy * dy) < g.obs[(int64_t)(3 * m + 2) * g.R + r] + g.d_safe_mpc);
    }
    if (hit) g.counters[3 * g.R + r] = 1;
}
#endif

int roll_fill(RollArgs& g, int64_t R, const b2c_rollout_params* p) {
    if (p->mode < B2C_ROLL_LQR || p->mode > B2C_ROLL_HYBRID) { b2c_set_error("bad rollout mode"); return -2; }
    if (p->ticks < 0 || p->table_rows < 1 || p->mpc.M < 0) { b2c_set_error("bad rollout sizes"); return -2; }
    if (p->mode != B2C_ROLL_LQR && (p->mpc.kind != B2C_MPC_LTV || p->mpc.N < 1)) {
        b2c_set_error("rollouts solve the solve_with_ltv problem (kind = B2C_MPC_LTV)"); return -2; }
    g.R = R; g.mode = p->mode; g.t = 0; g.mpc_rate = p->mpc_rate > 0 ? p->mpc_rate : 1;
    g.min_dwell = p->min_dwell; g.rows = p->table_rows; g.M = p->mpc.M; g.N = p->mpc.N;
    g.dt = p->dt; g.plant_v = p->plant_v_max; g.plant_w = p->plant_w_max; g.lqr_v = p->lqr_v_max; g.lqr_w = p->lqr_w_max;
    g.d_safe_risk = p->risk_d_safe; g.d_trigger = p->risk_d_trigger;
    g.alpha = p->risk_alpha;       // the NORMALISED weight RiskMetrics holds after __init__ (:76-79)
    g.threshold_low = p->risk_threshold_low; g.d_safe_mpc = p->mpc.d_safe;
    return 0;
}

}  // namespace

#ifndef B2C_HOST_EMUL
static size_t al256(size_t b) { return (b + 255) / 256 * 256; }

int b2c_rollout_dev(b2c_ctx* ctx, int64_t R, const b2c_rollout_params* p, const double* table,
                    const int32_t* k0, double* x, const double* obs, int32_t* mpc_step_count,
                    int32_t* counters, double* stats,
                    double* log_x, double* log_u, double* log_risk, int8_t* log_ctrl) {
    if (!ctx || !p || !table || !k0 || !x || !counters || !stats) { b2c_set_error("null argument"); return -1; }
    if (R <= 0) return 0;
    if (R >= ((int64_t)1 << 31)) { b2c_set_error("at most 2^31-1 robots per call"); return -2; }
    RollArgs g;
    int rc = roll_fill(g, R, p);
    if (rc) return rc;
    if (g.M > 0 && !obs) { b2c_set_error("obstacles missing"); return -1; }
    B2C_DEVICE_GUARD(ctx);
    const bool qp = p->mode != B2C_ROLL_LQR;
    const int M = g.M, rows = g.rows;
    // ---- working memory (context-owned arena, apart from the K3 scratch) ------------------------------------
    size_t need = al256(6 * rows * 8) + 2 * al256(rows * 8) + 2 * al256(rows * 4)     // gain table, v/th columns
                + al256(2 * R * 8) + 2 * al256(R * 4) + al256(R * 4) + 256              // u, prev, since, work, nwork
                + al256(R * 4);                                                         // step counters if not given
    if (qp) need += al256(3 * R * 8) + al256(R * 4) + al256((3 * M + 1) * R * 8) + al256(2 * R * 8) + al256(R * 8)
                  + 4 * al256(R * 4);
    rc = b2c_reserve(&ctx->dstage, &ctx->dstage_bytes, need, false);
    if (rc) return rc;
    char* base = (char*)ctx->dstage; size_t off = 0;
    auto take = [&](size_t bytes) { void* q = base + off; off += al256(bytes); return q; };
    double* ktab = (double*)take(6 * rows * 8);
    double* tv = (double*)take(rows * 8); double* tth = (double*)take(rows * 8);
    int32_t* kit = (int32_t*)take(rows * 4); int32_t* kst = (int32_t*)take(rows * 4);
    g.u = (double*)take(2 * R * 8); g.prev = (int32_t*)take(R * 4); g.since = (int32_t*)take(R * 4);
    g.work = (int32_t*)take(R * 4); g.nwork = (unsigned long long*)take(256);
    int32_t* own_sc = (int32_t*)take(R * 4);
    double *q_x0 = nullptr, *q_ob = nullptr, *q_u0 = nullptr, *q_cost = nullptr;
    int32_t *q_k0 = nullptr, *q_sc = nullptr, *q_st = nullptr, *q_it = nullptr, *q_sl = nullptr;
    if (qp) {
        q_x0 = (double*)take(3 * R * 8); q_k0 = (int32_t*)take(R * 4);
        q_ob = (double*)take((3 * M + 1) * R * 8); q_u0 = (double*)take(2 * R * 8); q_cost = (double*)take(R * 8);
        q_sc = (int32_t*)take(R * 4); q_st = (int32_t*)take(R * 4); q_it = (int32_t*)take(R * 4); q_sl = (int32_t*)take(R * 4);
    }
    cudaStream_t s = ctx->stream;
    int32_t* step_count = mpc_step_count;
    if (!step_count) {                                // fresh controllers: _step_count = 0 (mpc_controller.py:143)
        step_count = own_sc;
        B2C_CUDA(cudaMemsetAsync(own_sc, 0, R * 4, s));
    }
    g.table = table; g.ktab = ktab; g.k0 = k0; g.x = x; g.obs = obs; g.counters = counters; g.stats = stats;
    g.log_x = log_x; g.log_u = log_u; g.log_risk = log_risk; g.log_ctrl = log_ctrl;
    // ---- LQR gain per table row (K2, guard of compute_gain) ---------------------------------------------------
    if (p->mode != B2C_ROLL_MPC) {
        k_roll_table_cols<<<b2c_blocks(rows, 256), 256, 0, s>>>(rows, table, tv, tth);
        B2C_CHECK_LAUNCH(ctx);
        rc = b2c_dare_dev(ctx, rows, B2C_LIN_EULER, B2C_GUARD_LQR, p->dt, p->lqr_Qd, p->lqr_Rd, tv, tth, ktab, nullptr, kit, kst);
        if (rc) return rc;
    }
    k_roll_init<<<b2c_blocks(R, 256), 256, 0, s>>>(g);
    B2C_CHECK_LAUNCH(ctx);
    const int rblocks = b2c_blocks(R, 256);
    for (int t = 0; t < p->ticks; ++t) {
        g.t = t;
        const bool may_solve = qp && !(p->mode == B2C_ROLL_MPC && (t % g.mpc_rate) != 0);
        if (may_solve) B2C_CUDA(cudaMemsetAsync(g.nwork, 0, sizeof(unsigned long long), s));
        k_roll_decide<<<rblocks, 256, 0, s>>>(g);
        B2C_CHECK_LAUNCH(ctx);
        if (may_solve) {
            // the QP batch of this tick: R is the capacity (and the SoA stride), *g.nwork the number of instances
            k_roll_gather<<<rblocks, 256, 0, s>>>(g, R, q_x0, q_k0, q_ob, q_sc, step_count);
            B2C_CHECK_LAUNCH(ctx);
            MpcArgs a;
            mpc_fill_args(a, R, &p->mpc, q_x0, nullptr, nullptr, q_ob, q_sc, q_u0, q_cost, q_st, q_it, q_sl, nullptr, nullptr);
            a.n_dev = g.nwork;
            a.table = table; a.table_rows = rows; a.k0 = q_k0;
            rc = mpc_table_launch(ctx, a, p->mpc.engine);
            if (rc) return rc;
            k_roll_scatter<<<rblocks, 256, 0, s>>>(g, R, q_u0, q_sc, q_st, step_count);
            B2C_CHECK_LAUNCH(ctx);
        }
        k_roll_plant<<<rblocks, 256, 0, s>>>(g);
        B2C_CHECK_LAUNCH(ctx);
    }
    B2C_CUDA(cudaStreamSynchronize(s));
    return 0;
}

// host pointers, same [field][robot] layout as the device entry point (the Python mirror transposes with numpy)
int b2c_rollout_host(b2c_ctx* ctx, int64_t R, const b2c_rollout_params* p, const double* table,
                     const int32_t* k0, double* x, const double* obs, int32_t* mpc_step_count,
                     int32_t* counters, double* stats,
                     double* log_x, double* log_u, double* log_risk, int8_t* log_ctrl) {
    if (!ctx || !p || !table || !k0 || !x || !counters || !stats) { b2c_set_error("null argument"); return -1; }
    if (R <= 0) return 0;
    B2C_DEVICE_GUARD(ctx);
    const int M = p->mpc.M, T = p->ticks;
    const size_t b_tab = (size_t)p->table_rows * 6 * 8, b_k0 = R * 4, b_x = 3 * R * 8, b_obs = (size_t)(3 * M) * R * 8 + 8,
                 b_sc = R * 4, b_cnt = 5 * R * 4, b_st = 3 * R * 8, b_lx = log_x ? (size_t)(T + 1) * 3 * R * 8 : 0,
                 b_lu = log_u ? (size_t)T * 2 * R * 8 : 0, b_lr = log_risk ? (size_t)T * R * 8 : 0,
                 b_lc = log_ctrl ? (size_t)T * R : 0;
    const size_t total = al256(b_tab) + al256(b_k0) + al256(b_x) + al256(b_obs) + al256(b_sc) + al256(b_cnt) +
                         al256(b_st) + al256(b_lx) + al256(b_lu) + al256(b_lr) + al256(b_lc) + 256;
    int rc = b2c_reserve(&ctx->aux2, &ctx->aux2_bytes, total, false);       // context-owned: no cudaMalloc per call
    if (rc) return rc;
    char* d = (char*)ctx->aux2;
    size_t off = 0;
    auto take = [&](size_t bytes) { char* q = d + off; off += al256(bytes); return q; };
    double* d_tab = (double*)take(b_tab); int32_t* d_k0 = (int32_t*)take(b_k0); double* d_x = (double*)take(b_x);
    double* d_obs = (double*)take(b_obs); int32_t* d_sc = (int32_t*)take(b_sc); int32_t* d_cnt = (int32_t*)take(b_cnt);
    double* d_st = (double*)take(b_st);
    double* d_lx = log_x ? (double*)take(b_lx) : nullptr; double* d_lu = log_u ? (double*)take(b_lu) : nullptr;
    double* d_lr = log_risk ? (double*)take(b_lr) : nullptr; int8_t* d_lc = log_ctrl ? (int8_t*)take(b_lc) : nullptr;
    cudaStream_t s = ctx->stream;
    cudaError_t e = cudaSuccess;
    auto H2D = [&](void* dst, const void* src, size_t n) { if (e == cudaSuccess && n) e = cudaMemcpyAsync(dst, src, n, cudaMemcpyHostToDevice, s); };
    auto D2H = [&](void* dst, const void* src, size_t n) { if (e == cudaSuccess && n) e = cudaMemcpyAsync(dst, src, n, cudaMemcpyDeviceToHost, s); };
    H2D(d_tab, table, b_tab); H2D(d_k0, k0, b_k0); H2D(d_x, x, b_x);
    if (M > 0 && obs) H2D(d_obs, obs, (size_t)(3 * M) * R * 8);
    if (mpc_step_count) H2D(d_sc, mpc_step_count, b_sc);
    if (e == cudaSuccess)
        rc = b2c_rollout_dev(ctx, R, p, d_tab, d_k0, d_x, d_obs, mpc_step_count ? d_sc : nullptr, d_cnt, d_st,
                             d_lx, d_lu, d_lr, d_lc);
    if (rc == 0) {
        D2H(x, d_x, b_x); D2H(counters, d_cnt, b_cnt); D2H(stats, d_st, b_st);
        if (mpc_step_count) D2H(mpc_step_count, d_sc, b_sc);
        D2H(log_x, d_lx, b_lx); D2H(log_u, d_lu, b_lu); D2H(log_risk, d_lr, b_lr); D2H(log_ctrl, d_lc, b_lc);
    }
    const cudaError_t e2 = cudaStreamSynchronize(s);  // nothing targets the caller's buffers after the return
    if (rc) return rc;
    if (e != cudaSuccess) return b2c_cuda_fail(e, "rollout host copies", __FILE__, __LINE__);
    if (e2 != cudaSuccess) return b2c_cuda_fail(e2, "rollout", __FILE__, __LINE__);
    return 0;
}

// ---- synthetic robots of BASELINE config 5 / SURVEY 8(d), generated on the device --------------------------------
// Counter-based stream keyed by (seed, GLOBAL robot index): robot i is the same robot whatever the world size or the
// shard it lands in, and a 1,048,576-robot shard needs no host->device copy.  Draw j of robot i is splitmix64 of
// (seed ^ i * GOLDEN) + (j + 1) * GOLDEN, top 53 bits -> [0, 1).  Draws: 0 start row, 1-3 state offsets, 4 + 2m, 5 + 2m
// the shift of obstacle m.  hybrid_controller/workloads.config5_robots is the numpy mirror (bit-identical; tested).
struct GenArgs {
    int64_t R, first; unsigned long long seed;
    const double* table; int rows, k0_max, M;
    double pos_jitter, th_jitter, obs_jitter;
    double base[3 * B2C_GEN_MAX_OBS];
    int32_t* k0; double* x; double* obs;
};
__device__ __forceinline__ double gen_u01(unsigned long long state, int j) {
    unsigned long long z = state + (unsigned long long)(j + 1) * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    return (double)(z >> 11) * (1.0 / 9007199254740992.0);
}
__global__ void __launch_bounds__(256) k_roll_generate(const GenArgs g) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= g.R) return;
    const unsigned long long state = g.seed ^ ((unsigned long long)(g.first + r) * 0x9E3779B97F4A7C15ull);
    int k0 = (int)(gen_u01(state, 0) * (double)g.k0_max);
    if (k0 > g.k0_max - 1) k0 = g.k0_max - 1;
    g.k0[r] = k0;
    const double* row = g.table + (int64_t)(k0 < g.rows - 1 ? k0 : g.rows - 1) * 6;
    const double jit[3] = {g.pos_jitter, g.pos_jitter, g.th_jitter};
#pragma unroll
    for (int c = 0; c < 3; ++c)
        g.x[(int64_t)c * g.R + r] = row[1 + c] + B2C_MUL(2.0 * gen_u01(state, 1 + c) - 1.0, jit[c]);
    for (int m = 0; m < g.M; ++m) {
        g.obs[(int64_t)(3 * m) * g.R + r] = g.base[3 * m] + B2C_MUL(2.0 * gen_u01(state, 4 + 2 * m) - 1.0, g.obs_jitter);
        g.obs[(int64_t)(3 * m + 1) * g.R + r] = g.base[3 * m + 1] + B2C_MUL(2.0 * gen_u01(state, 5 + 2 * m) - 1.0, g.obs_jitter);
        g.obs[(int64_t)(3 * m + 2) * g.R + r] = g.base[3 * m + 2];
    }
}

int b2c_rollout_generate_dev(b2c_ctx* ctx, int64_t R, int64_t first_index, uint64_t seed,
                             const double* table, int32_t table_rows, int32_t k0_max,
                             const double* base_obstacles, int32_t M,
                             double pos_jitter, double th_jitter, double obs_jitter,
                             int32_t* k0, double* x, double* obs) {
    if (!ctx || !table || !k0 || !x || (M > 0 && (!base_obstacles || !obs))) { b2c_set_error("null argument"); return -1; }
    if (M < 0 || M > B2C_GEN_MAX_OBS || table_rows < 1 || k0_max < 1) { b2c_set_error("bad generator sizes"); return -2; }
    if (R <= 0) return 0;
    B2C_DEVICE_GUARD(ctx);
    GenArgs g;
    g.R = R; g.first = first_index; g.seed = seed; g.table = table; g.rows = table_rows; g.k0_max = k0_max; g.M = M;
    g.pos_jitter = pos_jitter; g.th_jitter = th_jitter; g.obs_jitter = obs_jitter;
    for (int e = 0; e < 3 * B2C_GEN_MAX_OBS; ++e) g.base[e] = e < 3 * M ? base_obstacles[e] : 0.0;
    g.k0 = k0; g.x = x; g.obs = obs;
    k_roll_generate<<<b2c_blocks(R, 256), 256, 0, ctx->stream>>>(g);
    B2C_CHECK_LAUNCH(ctx);
    return 0;
}
#else
// Test-only host emulation of the same per-robot code (gain table and QP solves through the emulated K2 / K3).
extern "C" int b2c_emul_dare(int64_t B, int method, int guard, double dt, const double Qd[3], const double Rd[2],
                             const double* v, const double* th, double* K, double* P, int32_t* iters, int32_t* status);
extern "C" int b2c_emul_mpc(int64_t B, const b2c_mpc_params* p, const double* x0, const double* xref,
                            const double* uref, const double* obs, int32_t* step_count, double* u0, double* cost,
                            int32_t* status, int32_t* iters, int32_t* slack_used, double* useq, double* xpred);
extern "C" int b2c_emul_rollout(int64_t R, const b2c_rollout_params* p, const double* table, const int32_t* k0,
                                double* x, const double* obs, int32_t* mpc_step_count, int32_t* counters,
                                double* stats, double* log_x, double* log_u, double* log_risk, int8_t* log_ctrl) {
    RollArgs g;
    int rc = roll_fill(g, R, p);
    if (rc) return rc;
    const int rows = g.rows, N = g.N, M = g.M;
    std::vector<double> ktab(6 * rows), tv(rows), tth(rows), u(2 * R, 0.0);
    std::vector<int32_t> kit(rows), kst(rows), prev(R, -1), since(R, 0), work(R), own_sc(R, 0);
    for (int k = 0; k < rows; ++k) { tv[k] = table[k * 6 + 4]; tth[k] = table[k * 6 + 3]; }
    if (p->mode != B2C_ROLL_MPC)
        b2c_emul_dare(rows, B2C_LIN_EULER, B2C_GUARD_LQR, p->dt, p->lqr_Qd, p->lqr_Rd, tv.data(), tth.data(),
                      ktab.data(), nullptr, kit.data(), kst.data());
    int32_t* step_count = mpc_step_count ? mpc_step_count : own_sc.data();
    g.table = table; g.ktab = ktab.data(); g.k0 = k0; g.x = x; g.obs = obs; g.counters = counters; g.stats = stats;
    g.u = u.data(); g.prev = prev.data(); g.since = since.data(); g.work = work.data(); g.nwork = nullptr;
    g.log_x = log_x; g.log_u = log_u; g.log_risk = log_risk; g.log_ctrl = log_ctrl;
    for (int64_t r = 0; r < R; ++r) {
        for (int c = 0; c < 5; ++c) counters[c * R + r] = 0;
        for (int c = 0; c < 3; ++c) stats[c * R + r] = 0.0;
        if (log_x) { log_x[r] = x[r]; log_x[R + r] = x[R + r]; log_x[2 * R + r] = x[2 * R + r]; }
        bool hit = false;
        for (int m = 0; m < M; ++m) {
            const double dx = x[r] - obs[(3 * m) * R + r], dy = x[R + r] - obs[(3 * m + 1) * R + r];
            hit = hit || (sqrt(dx * dx + dy * dy) < obs[(3 * m + 2) * R + r] + g.d_safe_mpc);
        }
        if (hit) counters[3 * R + r] = 1;
    }
    std::vector<double> q_x0(3 * R), q_xr(3 * (N + 1) * R), q_ur(2 * N * R + 1), q_ob(3 * M * R + 1), q_u0(2 * R), q_cost(R);
    std::vector<int32_t> q_sc(R), q_st(R), q_it(R), q_sl(R), q_k0(R);
    for (int t = 0; t < p->ticks; ++t) {
        g.t = t;
        int64_t n = 0;
        for (int64_t r = 0; r < R; ++r) if (roll_decide_one(g, r)) work[n++] = (int32_t)r;
        if (n > 0) {
            for (int64_t j = 0; j < n; ++j) {
                roll_gather_one(g, j, n, q_x0.data(), q_k0.data(), q_ob.data(), q_sc.data(), step_count);
                for (int k = 0; k <= N; ++k) {            // get_trajectory_segment(row, N+1), clamped at the end of the table
                    const int rr = q_k0[j] + k < rows - 1 ? q_k0[j] + k : rows - 1;
                    const double* row = table + (int64_t)rr * 6;
                    q_xr[(3 * k) * n + j] = row[1]; q_xr[(3 * k + 1) * n + j] = row[2]; q_xr[(3 * k + 2) * n + j] = row[3];
                    if (k < N) { q_ur[(2 * k) * n + j] = row[4]; q_ur[(2 * k + 1) * n + j] = row[5]; }
                }
            }
            b2c_emul_mpc(n, &p->mpc, q_x0.data(), q_xr.data(), q_ur.data(), q_ob.data(), q_sc.data(), q_u0.data(),
                         q_cost.data(), q_st.data(), q_it.data(), q_sl.data(), nullptr, nullptr);
            for (int64_t j = 0; j < n; ++j) roll_scatter_one(g, j, n, q_u0.data(), q_sc.data(), q_st.data(), step_count);
        }
        for (int64_t r = 0; r < R; ++r) roll_plant_one(g, r);
    }
    return 0;
}
#endif
