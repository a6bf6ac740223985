// Kernel-argument block and small helpers shared by the two K3 kernels (mpc.cu: general path with
// per-instance scratch in HBM; mpc_slots.cu: persistent slot kernel with the per-stage data in
// shared memory).  Not part of the C ABI.
#pragma once
#include "b2c_common.cuh"

struct MpcArgs {
    int64_t B;                   // instances (and the stride of every SoA buffer)
    const unsigned long long* n_dev;   // optional: the number of instances actually present (<= B), read on the device --
                                       // the rollout engine launches K3 on a work list whose length the host never sees
    int kind, N, M, bs, Nb, max_iter, ramp_up_steps, soft;
    double q0, q1, q2, r0, r1, p0, p1, p2;
    double d_safe, rho, v_max, w_max, dt, tol;
    const double* __restrict__ x0;
    const double* __restrict__ xref;
    const double* __restrict__ uref;
    const double* __restrict__ obs;
    int32_t* step_count;
    double* u0; double* cost; int32_t* status; int32_t* iters; int32_t* slack_used;
    double* useq; double* xpred;
    double* scratch;
    int lam_off;                 // general kernel, hard rows: first scratch field of the N*M multipliers
    // slot path only ------------------------------------------------------------------------------------------
    // input / output layouts.  in_aos: x0 (B,3), xref (B,N+1,3), uref (B,N,2), obs (B,M,3) as a numpy user holds them
    // (the *_host entry points hand the staged buffers over as they are: no re-layout kernels); 0: instance-minor SoA.
    // out_aos: u0 (B,2), useq (B,N,2), xpred (B,N+1,3).  table != NULL: (x_refs, u_refs) of instance i are rows
    // min(k0[i] + k, table_rows - 1) of the shared reference table (rows,6) [t,px,py,th,v,w], i.e.
    // ReferenceTrajectoryGenerator.get_trajectory_segment(k0[i], N+1) (HC/trajectory/reference_generator.py:299-326).
    int in_aos, out_aos;
    const double* __restrict__ table; int table_rows; const int32_t* __restrict__ k0;
    // prepared records (k_mpc_prepare -> k_mpc_slots): tiles of 32 instances,
    //   stage k of lane l : rec + tile*rec_tile + (k*32 + l)*6   {cos*dt, sin*dt, u_ref0, u_ref1, p_ref x, p_ref y}
    //   header of lane l  : rec + tile*rec_tile + 6*N*32 + l*rec_hdr   {z0x, z0y, z0t, -, (ox, oy, d_safe+r) x M}
    double* rec; int rec_hdr; int64_t rec_tile;
    unsigned long long* queue;   // next unclaimed instance index
    int64_t nslots;              // resident slots (= grid * block)
    int mir;                     // slot kernel: the gains mirror is held in tensor memory (set by the launcher)
    int fine_classes;            // slot kernel: six difficulty classes instead of four (set by the launcher)
    unsigned long long* class_count;   // instances per difficulty class (claim order), or unused
    uint32_t* perm;              // [class][B] instance ids in claim order; NULL = claim in index order
};

B2C_HD int64_t mpc_count(const MpcArgs& a) {
#ifdef __CUDA_ARCH__
    if (a.n_dev) { const int64_t n = (int64_t)*a.n_dev; return n < a.B ? n : a.B; }
#endif
    return a.B;
}

// input accessors of the slot path (layout flags are launch-uniform)
B2C_HD int mpc_table_row(const MpcArgs& a, int64_t i, int k) {
    const int r = a.k0[i] + k;
    return r < a.table_rows - 1 ? r : a.table_rows - 1;
}
B2C_HD double mpc_in_x0(const MpcArgs& a, int64_t i, int c) {
    return B2C_LDG(a.x0 + (a.in_aos ? i * 3 + c : (int64_t)c * a.B + i));
}
B2C_HD double mpc_in_xref(const MpcArgs& a, int64_t i, int k, int c) {
    if (a.table) return B2C_LDG(a.table + (int64_t)mpc_table_row(a, i, k) * 6 + 1 + c);
    return B2C_LDG(a.xref + (a.in_aos ? i * (3 * (a.N + 1)) + 3 * k + c : (int64_t)(3 * k + c) * a.B + i));
}
B2C_HD double mpc_in_uref(const MpcArgs& a, int64_t i, int k, int c) {
    if (a.table) return B2C_LDG(a.table + (int64_t)mpc_table_row(a, i, k) * 6 + 4 + c);
    return B2C_LDG(a.uref + (a.in_aos ? i * (2 * a.N) + 2 * k + c : (int64_t)(2 * k + c) * a.B + i));
}
B2C_HD double mpc_in_obs(const MpcArgs& a, int64_t i, int e) {
    return B2C_LDG(a.obs + (a.in_aos ? i * (3 * a.M) + e : (int64_t)e * a.B + i));
}
B2C_HD int64_t mpc_out_u(const MpcArgs& a, int64_t i, int k, int c) {       // useq element (k, c)
    return a.out_aos ? i * (2 * a.N) + 2 * k + c : (int64_t)(2 * k + c) * a.B + i;
}
B2C_HD int64_t mpc_out_x(const MpcArgs& a, int64_t i, int k, int c) {
    return a.out_aos ? i * (3 * (a.N + 1)) + 3 * k + c : (int64_t)(3 * k + c) * a.B + i;
}

// np.unwrap step (numpy/lib/function_base.py semantics, period 2pi, discont pi)
B2C_HD double unwrap_correction(double dd) {
    const double PI = 3.141592653589793, TWO_PI = 6.283185307179586;
    if (fabs(dd) < PI) return 0.0;            // the common case: no fmod
    double m = fmod(dd + PI, TWO_PI);
    if (m < 0.0) m += TWO_PI;                 // python-style modulo
    double ddmod = m - PI;
    if (ddmod == -PI && dd > 0.0) ddmod = PI;
    return ddmod - dd;
}

static inline void mpc_fill_args(MpcArgs& a, int64_t B, const b2c_mpc_params* p,
                                 const double* x0, const double* xref, const double* uref, const double* obs,
                                 int32_t* step_count, double* u0, double* cost, int32_t* status,
                                 int32_t* iters, int32_t* slack_used, double* useq, double* xpred) {
    a.B = B; a.n_dev = nullptr; a.kind = p->kind; a.N = p->N; a.M = p->M;
    a.bs = (p->kind == B2C_MPC_LTV && p->block_size > 1) ? p->block_size : 1;
    a.Nb = (a.N + a.bs - 1) / a.bs;
    a.max_iter = p->max_iter > 0 ? p->max_iter : ((p->soft || p->M == 0) ? 60 : 2000);   // hard rows: all multiplier rounds together
    a.ramp_up_steps = p->ramp_up_steps; a.soft = p->soft;
    a.q0 = p->Qd[0]; a.q1 = p->Qd[1]; a.q2 = p->Qd[2]; a.r0 = p->Rd[0]; a.r1 = p->Rd[1];
    a.p0 = p->Pd[0]; a.p1 = p->Pd[1]; a.p2 = p->Pd[2];
    a.d_safe = p->d_safe; a.rho = p->rho; a.v_max = p->v_max; a.w_max = p->w_max; a.dt = p->dt;
    a.tol = p->tol > 0 ? p->tol : 1e-14;
    a.x0 = x0; a.xref = xref; a.uref = uref; a.obs = obs; a.step_count = step_count;
    a.u0 = u0; a.cost = cost; a.status = status; a.iters = iters; a.slack_used = slack_used;
    a.useq = useq; a.xpred = xpred;
    a.scratch = nullptr; a.lam_off = 0; a.queue = nullptr; a.nslots = 0; a.mir = 0; a.fine_classes = 0; a.class_count = nullptr; a.perm = nullptr;
    a.in_aos = 0; a.out_aos = 0; a.table = nullptr; a.table_rows = 0; a.k0 = nullptr;
    a.rec = nullptr; a.rec_hdr = 0; a.rec_tile = 0;
}

// mpc_slots.cu
bool    mpc_slots_eligible(const MpcArgs& a, int* tpb_out);
int     mpc_slots_launch(b2c_ctx* ctx, MpcArgs& a, int tpb);
int64_t mpc_slots_scratch_bytes(const MpcArgs& a, int sm_count);
// mpc.cu: run (B, params) through the engine the parameters select; `a` carries pointers and layout flags
int     mpc_launch(b2c_ctx* ctx, MpcArgs& a, int engine);
int     mpc_table_launch(b2c_ctx* ctx, MpcArgs& a, int engine);   // a.table / a.k0 set; SoA x0, obs, outputs
bool    mpc_uses_slots(const MpcArgs& a, int engine);
#define MPC_LS_CLOSED 3    // closed-loop trials (alpha = 1, 1/2, 1/4) before the projection arc
#define MPC_LS_MAX 18
#define MPC_HARD_TOL 1e-9    // hard rows: KKT residual (violation / complementarity, metres) at which the multiplier loop stops
#define MPC_HARD_OUTER 120   // multiplier updates before the rows are declared infeasible
#define MPC_HARD_MU_MAX 1e6  // stiffest penalty tried (beyond it the piecewise-quadratic inner problem chatters)
// First trial step of a Newton iteration: min(1, MPC_LS_GROW * the step the previous iteration accepted).
// Full steps stay full (quadratic convergence is untouched: after an accepted alpha = 1 the next trial is 1);
// an instance that is crossing kinks of the obstacle penalty, where only alpha ~ 1/64 descends, no longer
// pays six rejected trials per iteration.  With the arc continuing from the closed-loop alpha instead of
// restarting at 1, the worst instance of config 4 needs 27 forward sweeps instead of 44 (same iterations).
#define MPC_LS_GROW 4.0
