/*
 * b2c.h -- C ABI of the B200 batched controller engine (libb2c.so).
 *
 * Drop-in boundary for the data-parallel hot path of the reference
 * (citations: /root/reference/src/hybrid_controller/hybrid_controller/ = HC/):
 *   b2c_linearize*   <- Linearizer.get_jacobians / discretize_euler / discretize_exact /
 *                       get_discrete_model / get_discrete_model_explicit
 *                       (HC/models/linearization.py:62-225)
 *   b2c_dare*        <- LQRController.compute_gain / get_lqr_gain
 *                       (HC/controllers/lqr_controller.py:92-147, :217-242)
 *   b2c_lqr_control* <- LQRController.compute_control (+ _normalize_angle, _clip_control)
 *                       (HC/controllers/lqr_controller.py:149-189, :244-257)
 *   b2c_mpc*         <- MPCController.solve_with_ltv (HC/controllers/mpc_controller.py:345-522),
 *                       MPCController.solve (:150-314), _get_fallback_solution (:316-343)
 *   b2c_rollout_dev  <- the closed-loop loops of run_simulation.py (:58-96, :225-280, :500-570) with
 *                       RiskMetrics.assess_risk and DifferentialDriveRobot.simulate_step
 *
 * Conventions
 *   - all arithmetic and all real buffers are IEEE fp64 (the reference is numpy fp64);
 *   - "_dev" entry points take DEVICE pointers in instance-minor SoA layout
 *     (element [f][i] at  f*B + i ; i = instance, coalesced across a warp);
 *   - "_host" entry points take HOST pointers in the natural instance-major layout a numpy
 *     user holds ((B,3,3), (B,N+1,3) ...): the library stages them through pinned memory,
 *     copies, re-lays them out on the device, runs the same kernels and copies results back;
 *   - caller allocates every in/out buffer; the library owns only the opaque context
 *     (stream, scratch, pinned staging);
 *   - return value: 0 = ok, <0 = error (b2c_last_error() gives the text). Per-instance solver
 *     outcomes are reported in status[] and never as an error code;
 *   - a context is not thread-safe: one per host thread / per GPU.  Every entry point runs on the context's device
 *     whatever the caller's current device is, and leaves the caller's current device unchanged;
 *   - there is NO CPU fallback: without a CUDA device b2c_create fails.
 */
#ifndef B2C_H
#define B2C_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b2c_ctx b2c_ctx;

/* linearization method ids (Linearizer.get_discrete_model 'method', linearization.py:160-188) */
enum { B2C_LIN_JACOBIAN = 0,  /* continuous (A,B)            :62-96   */
       B2C_LIN_EULER    = 1,  /* I + A dt, B dt              :98-118 (== explicit :190-225) */
       B2C_LIN_EXACT    = 2   /* expm ZOH, closed form (A^2=0) :120-158 */ };

/* low-speed guard applied to v before linearizing */
enum { B2C_GUARD_NONE = 0,    /* get_lqr_gain                lqr_controller.py:217-242 */
       B2C_GUARD_LQR  = 1,    /* |v|<1e-6 -> 0.01            lqr_controller.py:120-122 */
       B2C_GUARD_MPC  = 2     /* |v|<=0.01 -> 0.1            mpc_controller.py:186,425 */ };

/* per-instance status of the iterative kernels */
enum { B2C_OK = 0, B2C_MAXITER = 1, B2C_NONFINITE = 2, B2C_INFEASIBLE = 3 };

enum { B2C_MPC_ENGINE_AUTO = 0, B2C_MPC_ENGINE_GENERAL = 1, B2C_MPC_ENGINE_SLOTS = 2 };

/* MPC problem kinds */
enum { B2C_MPC_LTV = 0,       /* solve_with_ltv: error state, LTV, move blocking, theta unwrap */
       B2C_MPC_LTI = 1        /* solve: absolute state, one linearization, no unwrap           */ };

typedef struct {
    int32_t kind;             /* B2C_MPC_LTV | B2C_MPC_LTI                                     */
    int32_t N;                /* horizon                      (mpc_controller.py:111)           */
    int32_t M;                /* obstacles per instance (0 allowed)                             */
    int32_t block_size;       /* move blocking, LTV only      (:118-121)                        */
    int32_t soft;             /* use_soft_constraints (:152, :347).  0 = hard obstacle rows (:197-200, :260-270,
                                 :383-386, :464-468): solved by the method of multipliers in the general kernel;
                                 rows that cannot be met give status B2C_INFEASIBLE and the fallback outputs   */
    int32_t max_iter;         /* Newton iteration cap (<=0: default 60; 2000 with hard rows, where it
                                 covers all multiplier rounds together)                          */
    int32_t ramp_up_steps;    /* cold-start ramp length       (:144), LTV only                  */
    int32_t engine;           /* kernel choice: 0 = automatic, 1 = general kernel (HBM scratch), 2 = persistent
                                 slot kernel (shared-memory + tensor-memory scratch; LTV, block_size 1, soft rows)
                                 or error -5 */
    double  Qd[3], Rd[2], Pd[3];
    double  d_safe, rho, v_max, w_max, dt;
    double  tol;              /* relative Newton-decrement tolerance (<=0: default 1e-14)       */
} b2c_mpc_params;

/* ---- lifecycle ------------------------------------------------------------------------- */
int         b2c_create(b2c_ctx** out, int device);           /* device<0: current device       */
int         b2c_create_on_stream(b2c_ctx** out, int device, void* cuda_stream);
void        b2c_destroy(b2c_ctx* ctx);
const char* b2c_last_error(void);
int         b2c_version(void);
int         b2c_current_device(int* out);                     /* the calling thread's current CUDA device */
int         b2c_sync(b2c_ctx* ctx);                           /* cudaStreamSynchronize          */
void*       b2c_stream(b2c_ctx* ctx);                         /* the cudaStream_t in use        */
/* total kernel launches issued through this context since creation */
int64_t     b2c_launch_count(b2c_ctx* ctx);

/* pinned host memory for the *_host entry points (full PCIe rate; pageable memory also works) */
int         b2c_host_alloc(void** out, int64_t bytes);
int         b2c_host_free(void* p);

/* ---- K1: Jacobian + discretization ----------------------------------------------------- */
/* dev: v[B], th[B] -> Ad[9][B], Bd[6][B] (row-major 3x3 / 3x2 element index as field)        */
int b2c_linearize_dev(b2c_ctx*, int64_t B, int method, int guard, double dt,
                      const double* v, const double* th, double* Ad, double* Bd);
/* dev, packed: the same result in numpy layout Ad (B,3,3), Bd (B,3,2)                        */
int b2c_linearize_packed_dev(b2c_ctx*, int64_t B, int method, int guard, double dt,
                             const double* v, const double* th, double* Ad, double* Bd);
int b2c_linearize_host(b2c_ctx*, int64_t B, int method, int guard, double dt,
                       const double* v, const double* th, double* Ad /*(B,3,3)*/,
                       double* Bd /*(B,3,2)*/);

/* ---- K2: DARE + LQR gain ---------------------------------------------------------------- */
/* dev: v[B], th[B] -> K[6][B] (row-major 2x3), P[6][B] upper triangle (xx,xy,xt,yy,yt,tt) or
 * NULL, iters[B] or NULL, status[B] or NULL.  method/guard as above (the reference uses
 * EULER + GUARD_LQR in compute_gain, EULER + GUARD_NONE in get_lqr_gain).                    */
int b2c_dare_dev(b2c_ctx*, int64_t B, int method, int guard, double dt,
                 const double Qd[3], const double Rd[2],
                 const double* v, const double* th,
                 double* K, double* P, int32_t* iters, int32_t* status);
/* host: K (B,2,3), P (B,3,3) full symmetric or NULL                                          */
int b2c_dare_host(b2c_ctx*, int64_t B, int method, int guard, double dt,
                  const double Qd[3], const double Rd[2],
                  const double* v, const double* th,
                  double* K, double* P, int32_t* iters, int32_t* status);

/* ---- LQR control law --------------------------------------------------------------------- */
/* dev: x[3][B], xref[3][B], uref[2][B], K[6][B] -> u[2][B], err[3][B] (wrapped) or NULL       */
int b2c_lqr_control_dev(b2c_ctx*, int64_t B, double v_max, double w_max,
                        const double* x, const double* xref, const double* uref,
                        const double* K, double* u, double* err);

/* host: numpy layouts x, xref (B,3); uref (B,2); K (B,2,3) -> u (B,2), err (B,3) or NULL                       */
int b2c_lqr_control_host(b2c_ctx*, int64_t B, double v_max, double w_max,
                         const double* x, const double* xref, const double* uref,
                         const double* K, double* u, double* err);

/* ---- K3: MPC QP ---------------------------------------------------------------------------- */
/* bytes of device scratch b2c_mpc_dev needs for (B, params); the context grows its own scratch
 * on demand, this is informational (memory planning for 180 GB HBM).                          */
int64_t b2c_mpc_scratch_bytes(const b2c_mpc_params* p, int64_t B);

/* dev (SoA): x0[3][B], xref[(N+1)*3][B] (field = k*3+c), uref[N*2][B], obs[M*3][B]
 * (field = m*3+{x,y,r}), step_count[B] (in: successful solves so far; out: +1 on success; may be
 * NULL = ramp off).  Outputs: u0[2][B], cost[B], status[B], iters[B], slack_used[B];
 * useq[N*2][B] / xpred[(N+1)*3][B] optional (NULL to skip).
 * For kind=LTI xref/uref must already hold N+1 / N rows (the host mirror pads, :172-183).    */
int b2c_mpc_dev(b2c_ctx*, int64_t B, const b2c_mpc_params* p,
                const double* x0, const double* xref, const double* uref, const double* obs,
                int32_t* step_count,
                double* u0, double* cost, int32_t* status, int32_t* iters, int32_t* slack_used,
                double* useq, double* xpred);
/* host (instance-major): x0 (B,3), xref (B,N+1,3), uref (B,N,2), obs (B,M,3), step_count (B);
 * outputs u0 (B,2), cost (B), status/iters/slack_used (B), useq (B,N,2), xpred (B,N+1,3).    */
int b2c_mpc_host(b2c_ctx*, int64_t B, const b2c_mpc_params* p,
                 const double* x0, const double* xref, const double* uref, const double* obs,
                 int32_t* step_count,
                 double* u0, double* cost, int32_t* status, int32_t* iters, int32_t* slack_used,
                 double* useq, double* xpred);

/* asynchronous form of b2c_mpc_host: returns as soon as the copies and kernels are enqueued; the outputs are in the
 * caller's buffers after b2c_wait().  Staging is double-buffered: at most two calls are in flight per context (a third
 * waits for the first), so step s+1's host->device copies overlap step s's solve.  The caller keeps every buffer of a
 * call alive and unmodified until b2c_wait() (or until two later calls have been issued).                             */
int b2c_mpc_host_async(b2c_ctx*, int64_t B, const b2c_mpc_params* p,
                       const double* x0, const double* xref, const double* uref, const double* obs,
                       int32_t* step_count,
                       double* u0, double* cost, int32_t* status, int32_t* iters, int32_t* slack_used,
                       double* useq, double* xpred);
int b2c_wait(b2c_ctx*);                                       /* all asynchronous host calls have completed */

/* Table form of solve_with_ltv: every instance takes its references from ONE shared table (rows,6) row-major
 * [t,px,py,th,v,w] = ReferenceTrajectoryGenerator.generate() (HC/trajectory/reference_generator.py:196-230):
 *     (x_refs, u_refs) of instance i = get_trajectory_segment(k0[i], N+1)                  (:299-326)
 *                                    = rows min(k0[i] + k, rows-1), k = 0..N  (clamped at the end of the table),
 * which is what every caller of the reference does before solve_with_ltv (run_simulation.py:247, :392, :553).  The
 * segment is gathered on the device: 28 + 24 M bytes per instance cross PCIe instead of 24 + 40 N + 24 (N+1) + 24 M.
 * Results are bit-identical to b2c_mpc_dev / b2c_mpc_host on the materialised segments.  kind must be B2C_MPC_LTV.
 * dev: table, k0[B] device pointers; x0[3][B], obs[3M][B] and the outputs in the SoA layout of b2c_mpc_dev.
 * host: x0 (B,3), obs (B,M,3), outputs as b2c_mpc_host.                                                           */
int b2c_mpc_table_dev(b2c_ctx*, int64_t B, const b2c_mpc_params* p,
                      const double* table, int32_t table_rows, const int32_t* k0,
                      const double* x0, const double* obs, int32_t* step_count,
                      double* u0, double* cost, int32_t* status, int32_t* iters, int32_t* slack_used,
                      double* useq, double* xpred);
int b2c_mpc_table_host(b2c_ctx*, int64_t B, const b2c_mpc_params* p,
                       const double* table, int32_t table_rows, const int32_t* k0,
                       const double* x0, const double* obs, int32_t* step_count,
                       double* u0, double* cost, int32_t* status, int32_t* iters, int32_t* slack_used,
                       double* useq, double* xpred);
int b2c_mpc_table_host_async(b2c_ctx*, int64_t B, const b2c_mpc_params* p,
                             const double* table, int32_t table_rows, const int32_t* k0,
                             const double* x0, const double* obs, int32_t* step_count,
                             double* u0, double* cost, int32_t* status, int32_t* iters, int32_t* slack_used,
                             double* useq, double* xpred);

/* Allocation-free variant of b2c_mpc_dev.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments, layouts and results as b2c_mpc_dev run with engine = 1 (the general
 * kernel, every problem kind / move blocking / hard rows; p->engine is ignored), but the kernel's per-instance
 * scratch lives in a file-scope device array of 2,097,152 doubles (16 MiB, held by every GPU that loads the library
 * until the process exits) instead of the context's arena: the call performs no cudaMalloc / cudaFree.
 * Limit: B x scratch doubles per instance <= 2,097,152 -- b2c_mpc_static_max_instances(p) gives the largest B for a
 * shape (4,369 for N = 20 / M = 5 / block 1; 14,563 for N = 6); a larger B is rejected with error -7 and a message.
 * All calls on one device share the array: calls that overlap on the GPU (different streams or contexts) race, so
 * keep them on one stream.  The array keeps the previous call's contents; nothing needs clearing (the kernel writes
 * every field before it reads it).                                                                                  */
int64_t b2c_mpc_static_max_instances(const b2c_mpc_params* p);
int b2c_mpc_dev_static(b2c_ctx*, int64_t B, const b2c_mpc_params* p,
                       const double* x0, const double* xref, const double* uref, const double* obs,
                       int32_t* step_count,
                       double* u0, double* cost, int32_t* status, int32_t* iters, int32_t* slack_used,
                       double* useq, double* xpred);

/* ---- closed-loop rollouts (BASELINE config 5) ------------------------------ */
/* Replaces the reference's per-robot Python loops (run_simulation.py): run_lqr_simulation :58-96,
 * run_mpc_simulation :225-280, run_hybrid_simulation :500-570, with RiskMetrics.assess_risk
 * (HC/controllers/risk_metrics.py:84-222, distance risk; the hybrid loop passes no prediction) and
 * DifferentialDriveRobot.simulate_step (HC/models/differential_drive.py:138-172) on the device.    */
enum { B2C_ROLL_LQR = 0, B2C_ROLL_MPC = 1, B2C_ROLL_HYBRID = 2 };

typedef struct {
    int32_t mode;             /* B2C_ROLL_*                                                        */
    int32_t ticks;            /* plant steps to simulate                                           */
    int32_t mpc_rate;         /* MPC mode: solve every mpc_rate ticks and hold (run_simulation.py:243) */
    int32_t min_dwell;        /* hybrid: ticks a controller is kept after a switch (:520)         */
    int32_t table_rows;       /* rows of the reference table                                       */
    int32_t reserved;
    double  dt;
    double  plant_v_max, plant_w_max;                 /* DifferentialDriveRobot limits (:52, :160) */
    double  lqr_Qd[3], lqr_Rd[2], lqr_v_max, lqr_w_max;
    double  risk_d_safe, risk_d_trigger;
    double  risk_alpha, risk_beta;                    /* RiskMetrics.alpha / .beta AFTER its constructor normalised
                                                         them (:76-79); beta multiplies the predictive risk, which the
                                                         hybrid loop never supplies (always 0)                       */
    double  risk_threshold_low;
    b2c_mpc_params mpc;       /* the QP solved on MPC ticks (kind = B2C_MPC_LTV); mpc.M = obstacles per robot */
} b2c_rollout_params;

/* dev: table (rows,6) row-major [t,px,py,th,v,w] as ReferenceTrajectoryGenerator.generate() returns it;
 * k0[R] start row per robot (row index is clamped to the last row as get_reference_at_index does);
 * x[3][R] in: initial states, out: final states; obs[3M][R]; mpc_step_count[R] in/out or NULL (= fresh
 * controllers).  Outputs: counters[5][R] = lqr_steps, mpc_solves, switches, collision_states,
 * mpc_fallbacks; stats[3][R] = sum / max / last of the position tracking error norm.  Optional per-tick
 * logs (NULL to skip): log_x[(ticks+1)][3][R], log_u[ticks][2][R], log_risk[ticks][R],
 * log_ctrl[ticks][R] (0 = LQR, 1 = MPC).  Synchronous (returns when the rollout has finished).   */
int b2c_rollout_dev(b2c_ctx*, int64_t R, const b2c_rollout_params* p, const double* table,
                    const int32_t* k0, double* x, const double* obs, int32_t* mpc_step_count,
                    int32_t* counters, double* stats,
                    double* log_x, double* log_u, double* log_risk, int8_t* log_ctrl);

/* host pointers, same [field][robot] layouts */
int b2c_rollout_host(b2c_ctx*, int64_t R, const b2c_rollout_params* p, const double* table,
                     const int32_t* k0, double* x, const double* obs, int32_t* mpc_step_count,
                     int32_t* counters, double* stats,
                     double* log_x, double* log_u, double* log_risk, int8_t* log_ctrl);

/* Synthetic robots of BASELINE config 5 (Monte-Carlo closed-loop rollouts), generated on the device from a
 * counter-based stream keyed by (seed, GLOBAL robot index first_index + r): a robot is the same whatever the world size
 * or shard, and a shard needs no host->device copy.  Per robot: start row k0 ~ U{0..k0_max-1}; state = table row k0
 * [px,py,th] + U(+-pos_jitter, +-pos_jitter, +-th_jitter); obstacle m = base_obstacles[m] (HOST pointer, (M,3) [x,y,r],
 * M <= 8) with centre shifted by U(+-obs_jitter)^2.  Outputs in the layouts b2c_rollout_dev takes: k0[R], x[3][R],
 * obs[3M][R] (device pointers).  No counterpart in the reference (its runs are single-robot); the numpy mirror used by
 * the tests is hybrid_controller/workloads.config5_robots.                                                           */
int b2c_rollout_generate_dev(b2c_ctx*, int64_t R, int64_t first_index, uint64_t seed,
                             const double* table, int32_t table_rows, int32_t k0_max,
                             const double* base_obstacles, int32_t M,
                             double pos_jitter, double th_jitter, double obs_jitter,
                             int32_t* k0, double* x, double* obs);

/* ---- measurement helpers -------------------------------------------------------------------- */
/* register-resident FMA microbenchmark: returns achieved TFLOP/s (2 flop per FMA) for fp64
 * (bits=64) or fp32 (bits=32); used for the FP64 roofline denominator MEASURED_PEAKS lacks.   */
int b2c_measure_fma_peak(b2c_ctx*, int bits, double* tflops);

#ifdef __cplusplus
}
#endif
#endif /* B2C_H */
