// Kernel-argument block and small helpers shared by the two K3 kernels (mpc.cu: general path with
// per-instance scratch in HBM; mpc_slots.cu: persistent slot kernel with the per-stage data in
// shared memory).  Not part of the C ABI.
#pragma once
#include "b2c_common.cuh"

struct MpcArgs {
    int64_t B;
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
    // slot kernel only
    unsigned long long* queue;   // next unclaimed instance index
    int64_t nslots;              // resident slots (= grid * block)
    unsigned long long* class_count;   // instances per difficulty class (claim order), or unused
    uint32_t* perm;              // [class][B] instance ids in claim order; NULL = claim in index order
};

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
    a.B = B; a.kind = p->kind; a.N = p->N; a.M = p->M;
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
    a.scratch = nullptr; a.lam_off = 0; a.queue = nullptr; a.nslots = 0; a.class_count = nullptr; a.perm = nullptr;
}

// mpc_slots.cu
bool mpc_slots_eligible(const MpcArgs& a, int* tpb_out);
int  mpc_slots_launch(b2c_ctx* ctx, MpcArgs& a, int tpb);
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
