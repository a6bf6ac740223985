// K3 (fast path): persistent "slot" kernel for solve_with_ltv-shaped batches with block_size == 1
// (HC/controllers/mpc_controller.py:345-522).  Same method as mpc.cu (slack elimination + projected
// semismooth Newton with a Riccati inner solve), different machine mapping:
//
//   * one CTA per SM, one lane = one SLOT; a slot owns a private strip of SHARED memory holding the
//     per-stage data of the instance it is working on:
//        W[2N]  current inputs            U[8N]  union { Riccati gains K(6) k(2) | stage model(6) + previous inputs(2) }
//        OB[3M] the instance's obstacles                         (10N + 3M doubles; 1,720 B at N=20, M=5)
//     laid out [field][lane], i.e. bank-conflict free.  Forward and backward sweeps never touch HBM for
//     the sequential part of the recursion.
//   * read-only per-stage constants (cos*dt, sin*dt, u_ref, p_ref) are written once per instance into a
//     slot-indexed global strip G[6N][nslots] (coalesced, L2-resident: 148*128 slots * 960 B = 18 MB) and
//     software-prefetched one stage ahead; the gains are mirrored to GG[8N][nslots] so that a rejected
//     line-search trial can be undone without re-running the Riccati sweep.
//   * LANE REFILL: a lane whose instance has converged writes its outputs and immediately claims the next
//     unsolved instance from a global queue (warp-aggregated atomicAdd), so a warp never idles on its
//     slowest instance.  Every round of the warp loop is  [backward | finish+refill+setup | forward].
//
// The per-slot functions are B2C_HD so tests/ can run the identical state machine on the host
// (-DB2C_HOST_EMUL, test-only library).
#include "mpc_args.cuh"

namespace {

struct SlotMem {
    double* w;  double* u;  double* ob;               // shared strips, element (f) at f*SS (host emulation: SS = 1)
    double* ring;                                     // this lane's cell of ring slot 0 (device only)
    double* ring4;                                    // the same memory as a three-stage ring of four (backward sweep)
    double* g;  double* gg; int64_t gs;               // global strips: stage k of this slot at g + k*6*gs / gg + k*8*gs
};
// 16-byte vector access to the global strips (6 resp. 8 contiguous doubles per stage and slot)
struct Stage6 { double ck, sk, uo0, uo1, prx, pry; };
B2C_HD Stage6 g_load(const SlotMem& m, int k) {
    const double2* p = reinterpret_cast<const double2*>(m.g + (int64_t)k * 6 * m.gs);
    const double2 a = p[0], b = p[1], c = p[2];
    return Stage6{a.x, a.y, b.x, b.y, c.x, c.y};
}
B2C_HD void g_load4(const SlotMem& m, int k, double& ck, double& sk, double& uo0, double& uo1) {
    const double2* p = reinterpret_cast<const double2*>(m.g + (int64_t)k * 6 * m.gs);
    const double2 a = p[0], b = p[1];
    ck = a.x; sk = a.y; uo0 = b.x; uo1 = b.y;
}
B2C_HD void g_store(const SlotMem& m, int k, double ck, double sk, double uo0, double uo1, double prx, double pry) {
    double2* p = reinterpret_cast<double2*>(m.g + (int64_t)k * 6 * m.gs);
    p[0] = make_double2(ck, sk); p[1] = make_double2(uo0, uo1); p[2] = make_double2(prx, pry);
}
B2C_HD void gg_store(const SlotMem& m, int j, double K00, double K01, double K02, double K10, double K11,
                     double K12, double k0, double k1) {
    double2* p = reinterpret_cast<double2*>(m.gg + (int64_t)j * 8 * m.gs);
    p[0] = make_double2(K00, K01); p[1] = make_double2(K02, K10); p[2] = make_double2(K11, K12);
    p[3] = make_double2(k0, k1);
}
#define MW(f)  m.w[(f) * SS]
#define MU(f)  m.u[(f) * SS]
#define MO(f)  m.ob[(f) * SS]

// Two-stage ring in shared memory for the read-only stage constants: stage k+2 is fetched with
// cp.async (LDGSTS, no register dependency) while stages k and k+1 are being computed, so the L2
// latency of the G strip never reaches the recursion.  Cell of lane l, ring slot r, in the warp's block: [(r*32 + l)*6 .. +6).
// `dep` is a value computed from the cell being overwritten: it orders the refill after the read.
template <int SS>
B2C_HD void ring_issue(const SlotMem& m, int k, bool valid, double dep) {
#ifdef __CUDA_ARCH__
    if (valid) {
        const double* src = m.g + (int64_t)k * 6 * m.gs;
        const unsigned dst = (unsigned)__cvta_generic_to_shared(m.ring + (size_t)(k & 1) * 6 * 32);
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n\t"
                     "cp.async.cg.shared.global [%0+16], [%1+16], 16;\n\t"
                     "cp.async.cg.shared.global [%0+32], [%1+32], 16;  // after %2"
                     :: "r"(dst), "l"(src), "d"(dep) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
#else
    (void)m; (void)k; (void)valid; (void)dep;
#endif
}
template <int SS>
B2C_HD Stage6 ring_take(const SlotMem& m, int k) {
#ifdef __CUDA_ARCH__
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    const double2* p = reinterpret_cast<const double2*>(m.ring + (size_t)(k & 1) * 6 * 32);
    const double2 a = p[0], b = p[1], c = p[2];
    return Stage6{a.x, a.y, b.x, b.y, c.x, c.y};
#else
    return g_load(m, k);
#endif
}

// The backward sweep needs four of the six constants (cos*dt, sin*dt, u_ref) and its stages are the short ones
// (600 cycles against an L2 round trip that can exceed 1,200 under load): the same 12 doubles per slot then hold a
// THREE-stage ring of four, stage j-3 fetched while j .. j-2 are computed.  Cell of lane l, ring slot r, in the
// warp's block: [(r*32 + l)*4 .. +4).
template <int SS>
B2C_HD void ringb_issue(const SlotMem& m, int k, int r, bool valid, double dep) {
#ifdef __CUDA_ARCH__
    if (valid) {
        const double* src = m.g + (int64_t)k * 6 * m.gs;
        const unsigned dst = (unsigned)__cvta_generic_to_shared(m.ring4 + (size_t)r * 4 * 32);
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n\t"
                     "cp.async.cg.shared.global [%0+16], [%1+16], 16;  // after %2"
                     :: "r"(dst), "l"(src), "d"(dep) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
#else
    (void)m; (void)k; (void)r; (void)valid; (void)dep;
#endif
}
template <int SS>
B2C_HD void ringb_take(const SlotMem& m, int k, int r, double& ck, double& sk, double& uo0, double& uo1) {
#ifdef __CUDA_ARCH__
    asm volatile("cp.async.wait_group 2;" ::: "memory");
    const double2* p = reinterpret_cast<const double2*>(m.ring4 + (size_t)r * 4 * 32);
    const double2 a = p[0], b = p[1];
    ck = a.x; sk = a.y; uo0 = b.x; uo1 = b.y;
#else
    (void)r;
    g_load4(m, k, ck, sk, uo0, uo1);
#endif
}

// shared-memory doubles per slot: strips W, U, OB (+ 12 for the two ring cells)
__host__ __device__ static inline int slot_fields(const MpcArgs& a) { return 10 * a.N + 3 * a.M; }

enum { SLOT_IDLE = 0, SLOT_BWD = 1, SLOT_FWD_GAIN = 2, SLOT_FWD_PLAIN = 3 };

struct SlotState {
    int64_t inst;
    int mode, it, ls, st;
    double alpha, J, dec, z0x, z0y, z0t, lamx, lamy, lamt, max_viol;
    double x0x, x0t, r0x, r0t;                        // kept for the fallback law
    double ur00, ur01;                                // first reference input (outputs)
    int rej;                                          // a trial step of this solve has been rejected (kernel: asks for helper lanes)
};

// ----- setup: claim instance i, stage its constants, w = clip(0) ---------------------------------
// Three steps so that a warp can share the work of the lanes that refill in a round (kernel: the
// (lane, stage) pairs of steps 1 and 3 are dealt out over all 32 lanes; host emulation: plain loops):
//   1. setup_load_stage : raw rows of stage k (theta_ref, u_ref, p_ref) -> the (still unused) U strip
//   2. setup_owner      : x0 / first reference row / obstacles, np.unwrap along the horizon (sequential,
//                         cheap: one compare per stage unless the heading crosses +-pi)
//   3. setup_stage      : sincos of the unwrapped heading -> G strip, w_k = clip(0) -> W strip
template <int SS>
B2C_HD void setup_load_stage(const MpcArgs& a, const SlotMem& m, int64_t i, int k) {
    const int64_t B = a.B;
    MU(8 * k + 0) = B2C_LDG(a.xref + (int64_t)(3 * k + 2) * B + i);
    MU(8 * k + 1) = B2C_LDG(a.uref + (int64_t)(2 * k) * B + i);
    MU(8 * k + 2) = B2C_LDG(a.uref + (int64_t)(2 * k + 1) * B + i);
    MU(8 * k + 3) = B2C_LDG(a.xref + (int64_t)(3 * k) * B + i);
    MU(8 * k + 4) = B2C_LDG(a.xref + (int64_t)(3 * k + 1) * B + i);
}

#ifndef B2C_HOST_EMUL
// the same five rows with cp.async (global -> shared without a register round trip): a warp puts the rows
// of every refilling lane in flight at once and waits for all of them together
template <int SS>
__device__ __forceinline__ void setup_load_stage_async(const MpcArgs& a, const SlotMem& m, int64_t i, int k) {
#ifdef __CUDA_ARCH__
    const int64_t B = a.B;
    const unsigned dst = (unsigned)__cvta_generic_to_shared(&MU(8 * k));
    const double* xr = a.xref + (int64_t)(3 * k) * B + i;
    const double* ur = a.uref + (int64_t)(2 * k) * B + i;
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" :: "r"(dst), "l"(xr + 2 * B) : "memory");
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" :: "r"(dst + 1 * SS * 8), "l"(ur) : "memory");
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" :: "r"(dst + 2 * SS * 8), "l"(ur + B) : "memory");
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" :: "r"(dst + 3 * SS * 8), "l"(xr) : "memory");
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" :: "r"(dst + 4 * SS * 8), "l"(xr + B) : "memory");
#endif
}
#endif

template <int SS, int MT>
B2C_HD void setup_owner(const MpcArgs& a, const SlotMem& m, SlotState& s, int64_t i) {
    const int N = a.N, M = (MT >= 0) ? MT : a.M;
    const int64_t B = a.B;
    s.inst = i; s.it = 0; s.ls = 0; s.rej = 0; s.st = B2C_MAXITER; s.alpha = 0.0; s.dec = 0.0; s.max_viol = 0.0;
    const double x0x = B2C_LDG(a.x0 + i), x0y = B2C_LDG(a.x0 + B + i), x0t = B2C_LDG(a.x0 + 2 * B + i);
    const double r0x = B2C_LDG(a.xref + i), r0y = B2C_LDG(a.xref + B + i), r0t = B2C_LDG(a.xref + 2 * B + i);
    if (MT > 0) {                                 // obstacle strip: x, y, d_safe + r -- all loads in flight, then the stores
        constexpr int M3 = (MT > 0) ? 3 * MT : 3;
        double ob[M3];
#pragma unroll
        for (int e = 0; e < M3; ++e) ob[e] = B2C_LDG(a.obs + (int64_t)e * B + i);
#pragma unroll
        for (int e = 0; e < M3; ++e) MO(e) = (e % 3 == 2) ? a.d_safe + ob[e] : ob[e];
    } else {
        for (int e = 0; e < M; ++e) {
            MO(3 * e)     = B2C_LDG(a.obs + (int64_t)(3 * e) * B + i);
            MO(3 * e + 1) = B2C_LDG(a.obs + (int64_t)(3 * e + 1) * B + i);
            MO(3 * e + 2) = a.d_safe + B2C_LDG(a.obs + (int64_t)(3 * e + 2) * B + i);
        }
    }
    s.x0x = x0x; s.x0t = x0t; s.r0x = r0x; s.r0t = r0t;
    s.z0x = x0x - r0x; s.z0y = x0y - r0y; s.z0t = b2c_wrap(x0t - r0t);       // :392-401, :421
    s.mode = SLOT_FWD_PLAIN;
}

// np.unwrap along the horizon, in place on the staged headings (after the stage rows have landed)
template <int SS>
B2C_HD void setup_unwrap(const MpcArgs& a, const SlotMem& m, SlotState& s, const bool any_jump) {
    const int N = a.N;
    s.ur00 = MU(1); s.ur01 = MU(2);
    if (!any_jump) return;                            // every |step| < pi: np.unwrap is the identity
    double th_prev = s.r0t, cum = 0.0;
    for (int k = 1; k < N; ++k) {
        const double th_raw = MU(8 * k + 0);
        cum += unwrap_correction(th_raw - th_prev);
        th_prev = th_raw;
        if (cum != 0.0) MU(8 * k + 0) = th_raw + cum;
    }
}

template <int SS>
B2C_HD void setup_stage(const MpcArgs& a, const SlotMem& m, int k) {
    const double th = MU(8 * k + 0), uo0 = MU(8 * k + 1), uo1 = MU(8 * k + 2);
    double sn, cs; sincos(th, &sn, &cs);
    g_store(m, k, cs * a.dt, sn * a.dt, uo0, uo1, MU(8 * k + 3), MU(8 * k + 4));
    MW(2 * k)     = b2c_clip(0.0, -a.v_max - uo0, a.v_max - uo0);
    MW(2 * k + 1) = b2c_clip(0.0, -a.w_max - uo1, a.w_max - uo1);
}

// the three steps for one lane on its own (host emulation, and the reference order of operations)
template <int SS, int MT>
B2C_HD void slot_setup(const MpcArgs& a, const SlotMem& m, SlotState& s, int64_t i) {
    for (int k = 0; k < a.N; ++k) setup_load_stage<SS>(a, m, i, k);
    setup_owner<SS, MT>(a, m, s, i);
    setup_unwrap<SS>(a, m, s, true);
    for (int k = 0; k < a.N; ++k) setup_stage<SS>(a, m, k);
}

// one soft obstacle row, exact arithmetic (:439-468): accumulates cost / gradient / Hessian if violated
B2C_HD void slot_row(const MpcArgs& a, double ddx, double ddy, double d2, double c, double zx, double zy,
                     double& J, double& mv, double& gx, double& gy, double& wxx, double& wxy, double& wyy) {
    const double inv = B2C_RSQRT(d2);
    const double nx = ddx * inv, ny = ddy * inv;
    const double viol = c - (d2 * inv + nx * zx + ny * zy);
    if (viol > 0.0) {
        mv = fmax(mv, viol);
        J += a.rho * viol * viol;
        const double rn0 = a.rho * nx, rn1 = a.rho * ny;
        wxx += rn0 * nx; wxy += rn0 * ny; wyy += rn1 * ny;
        gx -= rn0 * viol; gy -= rn1 * viol;
    }
}

// ----- forward sweep -------------------------------------------------------------------------------
// gain == false: evaluate cost + stage model at the current inputs.
// gain == true : w+ = clip(w - alpha*kf - K (z+ - z)), rolled out closed loop; W is updated IN PLACE and
//                the previous inputs are parked in U[8k+6..7] (the model overwrites the gains it just used).
// Obstacle rows: row (k,m) is violated iff  c*dist > q  with  q = d2 + dd.dz  (both sides of
// c - (dist + n.dz) > 0 multiplied by dist), i.e. iff  c|c|*d2 > q|q| -- no square root.  All M tests of a
// stage are independent (they pipeline); only rows that pass take the rsqrt path.
template <int SS, int MT>
B2C_HD double slot_forward(const MpcArgs& a, const SlotMem& m, const SlotState& s, const bool gain,
                           double& max_viol, double& lamx, double& lamy, double& lamt) {
    const int N = a.N, M = (MT >= 0) ? MT : a.M;
    constexpr int MR = (MT > 0) ? MT : 1;
    double OX[MR], OY[MR], OC[MR], OCC[MR];
    if (MT > 0) {
#pragma unroll
        for (int mm = 0; mm < MR; ++mm) {
            OX[mm] = MO(3 * mm); OY[mm] = MO(3 * mm + 1); OC[mm] = MO(3 * mm + 2);
            OCC[mm] = OC[mm] * fabs(OC[mm]);
        }
    }
    double zx = s.z0x, zy = s.z0y, zt = s.z0t;        // new trajectory
    double ox = zx, oy = zy, ot = zt;                 // old trajectory
    double J = 0.0, mv = 0.0;
    const double alpha = gain ? s.alpha : 0.0;
    ring_issue<SS>(m, 0, true, 0.0);
    ring_issue<SS>(m, 1, N > 1, 0.0);
    for (int k = 0; k < N; ++k) {
        const Stage6 cur = ring_take<SS>(m, k);
        const double ck = cur.ck, sk = cur.sk, uo0 = cur.uo0, uo1 = cur.uo1, prx = cur.prx, pry = cur.pry;
        ring_issue<SS>(m, k + 2, k + 2 < N, ck + pry);       // constants two stages ahead
        const double wo0 = MW(2 * k), wo1 = MW(2 * k + 1);
        double w0 = wo0, w1 = wo1;
        if (gain) {
            const double dx = zx - ox, dy = zy - oy, dth = zt - ot;
            w0 = wo0 - alpha * MU(8 * k + 6) - (MU(8 * k + 0) * dx + MU(8 * k + 1) * dy + MU(8 * k + 2) * dth);
            w1 = wo1 - alpha * MU(8 * k + 7) - (MU(8 * k + 3) * dx + MU(8 * k + 4) * dy + MU(8 * k + 5) * dth);
            w0 = b2c_clip(w0, -a.v_max - uo0, a.v_max - uo0);
            w1 = b2c_clip(w1, -a.w_max - uo1, a.w_max - uo1);
            MW(2 * k) = w0; MW(2 * k + 1) = w1;
        }
        double wxx = a.q0, wxy = 0.0, wyy = a.q1;
        double gx = a.q0 * zx, gy = a.q1 * zy;
        const double gt = a.q2 * zt;
        J += a.q0 * zx * zx + a.q1 * zy * zy + a.q2 * zt * zt;
        if (MT > 0) {
            unsigned hit = 0;
#pragma unroll
            for (int mm = 0; mm < MR; ++mm) {
                const double ddx = prx - OX[mm], ddy = pry - OY[mm];
                const double d2 = ddx * ddx + ddy * ddy;
                const double q = d2 + ddx * zx + ddy * zy;
                if ((d2 > 1e-4) && (OCC[mm] * d2 > q * fabs(q))) hit |= (1u << mm);
            }
            while (hit) {                                 // rare; lanes with different rows share an iteration
#ifdef __CUDA_ARCH__
                const int mm = __ffs(hit) - 1;
#else
                int mm = 0; while (!((hit >> mm) & 1u)) ++mm;
#endif
                hit &= hit - 1u;
                const double ddx = prx - MO(3 * mm), ddy = pry - MO(3 * mm + 1);
                slot_row(a, ddx, ddy, ddx * ddx + ddy * ddy, MO(3 * mm + 2), zx, zy, J, mv, gx, gy, wxx, wxy, wyy);
            }
        } else if (MT < 0) {
            for (int mm = 0; mm < M; ++mm) {
                const double ddx = prx - MO(3 * mm), ddy = pry - MO(3 * mm + 1), c = MO(3 * mm + 2);
                const double d2 = ddx * ddx + ddy * ddy;
                const double q = d2 + ddx * zx + ddy * zy;
                if ((d2 > 1e-4) && (c * fabs(c) * d2 > q * fabs(q)))
                    slot_row(a, ddx, ddy, d2, c, zx, zy, J, mv, gx, gy, wxx, wxy, wyy);
            }
        }
        const double ut0 = w0 + uo0, ut1 = w1 + uo1;
        J += a.r0 * ut0 * ut0 + a.r1 * ut1 * ut1;
        MU(8 * k + 0) = gx;  MU(8 * k + 1) = gy;  MU(8 * k + 2) = gt;
        MU(8 * k + 3) = wxx; MU(8 * k + 4) = wxy; MU(8 * k + 5) = wyy;
        MU(8 * k + 6) = wo0; MU(8 * k + 7) = wo1;
        const double vr = b2c_guard_v(uo0, B2C_GUARD_MPC);                       // :425
        const double a_k = -vr * sk, b_k = vr * ck;
        const double nzx = zx + a_k * zt + ck * w0, nzy = zy + b_k * zt + sk * w0;
        const double nox = ox + a_k * ot + ck * wo0, noy = oy + b_k * ot + sk * wo0;
        zt = zt + a.dt * w1; zx = nzx; zy = nzy;
        ot = ot + a.dt * wo1; ox = nox; oy = noy;
    }
    J += a.p0 * zx * zx + a.p1 * zy * zy + a.p2 * zt * zt;                       // terminal cost, :412
    lamx = a.p0 * zx; lamy = a.p1 * zy; lamt = a.p2 * zt;
    max_viol = mv;
    return J;
}

// ----- backward sweep ------------------------------------------------------------------------------
// Riccati recursion on the stored stage models with the block_size == 1 structure written out
// (A = I + [0 0 al; 0 0 be; 0 0 0], B = [ck 0; sk 0; 0 dt], Hzz = [wxx wxy 0; wxy wyy 0; 0 0 q2], Huu = R).
// Gains go to U (over the model they were computed from) and are mirrored to GG.
template <int SS, int MT>
B2C_HD double slot_backward(const MpcArgs& a, const SlotMem& m, const SlotState& s) {
    const int N = a.N;
    const double dt = a.dt;
    double s00 = a.p0, s01 = 0.0, s02 = 0.0, s11 = a.p1, s12 = 0.0, s22 = a.p2;
    double v0 = s.lamx, v1 = s.lamy, v2 = s.lamt;     // value-function gradient
    double l0 = v0, l1 = v1, l2 = v2;                 // true costate
    double dec = 0.0;
    // the mirror store of stage j+1 is issued one stage late, from registers nothing else wants in between
    // (issued in place, the next writes to its source registers stalled on the store queue)
    double p0 = 0.0, p1 = 0.0, p2 = 0.0, p3 = 0.0, p4 = 0.0, p5 = 0.0, p6 = 0.0, p7 = 0.0;
    ringb_issue<SS>(m, N - 1, 0, true, 0.0);
    ringb_issue<SS>(m, N - 2, 1, N > 1, 0.0);
    ringb_issue<SS>(m, N - 3, 2, N > 2, 0.0);
    int rs = 0;                                       // ring slot of stage j
    for (int j = N - 1; j >= 0; --j) {
        if (j < N - 1) gg_store(m, j + 1, p0, p1, p2, p3, p4, p5, p6, p7);
        double ck, sk, uo0, uo1;
        ringb_take<SS>(m, j, rs, ck, sk, uo0, uo1);
        ringb_issue<SS>(m, j - 3, rs, j >= 3, ck + uo1);     // constants three stages ahead, into the cell just read
        rs = (rs == 2) ? 0 : rs + 1;
        const double gz0 = MU(8 * j + 0), gz1 = MU(8 * j + 1), gz2 = MU(8 * j + 2);
        const double H0 = MU(8 * j + 3), H1 = MU(8 * j + 4), H3 = MU(8 * j + 5);
        const double w0 = MW(2 * j), w1 = MW(2 * j + 1);
        const double gu0 = a.r0 * (w0 + uo0), gu1 = a.r1 * (w1 + uo1);
        const double vr = b2c_guard_v(uo0, B2C_GUARD_MPC);
        const double al = -vr * sk, be = vr * ck;
        // true gradient wrt w_j decides which bounded inputs are frozen; then the costate update
        const double gt0 = gu0 + ck * l0 + sk * l1;
        const double gt1 = gu1 + dt * l2;
        {
            const double nl2 = gz2 + al * l0 + be * l1 + l2;
            l0 = gz0 + l0; l1 = gz1 + l1; l2 = nl2;
        }
        const double lb0 = -a.v_max - uo0, ub0 = a.v_max - uo0, lb1 = -a.w_max - uo1, ub1 = a.w_max - uo1;
        const bool c0 = (w0 <= lb0 && gt0 > 0.0) || (w0 >= ub0 && gt0 < 0.0);
        const bool c1 = (w1 <= lb1 && gt1 > 0.0) || (w1 >= ub1 && gt1 < 0.0);
        // third column of S*A
        const double t0 = al * s00 + be * s01 + s02;
        const double t1 = al * s01 + be * s11 + s12;
        const double t2 = al * s02 + be * s12 + s22;
        const double Q00 = H0 + s00, Q01 = H1 + s01, Q02 = t0;
        const double Q11 = H3 + s11, Q12 = t1;
        const double Q22 = a.q2 + al * t0 + be * t1 + t2;
        double U00 = ck * s00 + sk * s01, U01 = ck * s01 + sk * s11, U02 = ck * t0 + sk * t1;
        double U10 = dt * s02, U11 = dt * s12, U12 = dt * t2;
        double R00 = a.r0 + ck * U00 + sk * U01;
        double R01 = ck * U10 + sk * U11;
        double R11 = a.r1 + dt * dt * s22;
        const double qz0 = gz0 + v0, qz1 = gz1 + v1, qz2 = gz2 + al * v0 + be * v1 + v2;
        double qu0 = gu0 + ck * v0 + sk * v1;
        double qu1 = gu1 + dt * v2;
        // frozen inputs: selects, not branches (lanes of a warp freeze different inputs at every stage)
        R00 = c0 ? 1.0 : R00; U00 = c0 ? 0.0 : U00; U01 = c0 ? 0.0 : U01; U02 = c0 ? 0.0 : U02; qu0 = c0 ? 0.0 : qu0;
        R11 = c1 ? 1.0 : R11; U10 = c1 ? 0.0 : U10; U11 = c1 ? 0.0 : U11; U12 = c1 ? 0.0 : U12; qu1 = c1 ? 0.0 : qu1;
        R01 = (c0 || c1) ? 0.0 : R01;
        const double idet = B2C_RCP(R00 * R11 - R01 * R01);
        const double i00 = R11 * idet, i01 = -R01 * idet, i11 = R00 * idet;
        const double K00 = i00 * U00 + i01 * U10, K01 = i00 * U01 + i01 * U11, K02 = i00 * U02 + i01 * U12;
        const double K10 = i01 * U00 + i11 * U10, K11 = i01 * U01 + i11 * U11, K12 = i01 * U02 + i11 * U12;
        const double k0 = i00 * qu0 + i01 * qu1, k1 = i01 * qu0 + i11 * qu1;
        MU(8 * j + 0) = K00; MU(8 * j + 1) = K01; MU(8 * j + 2) = K02;
        MU(8 * j + 3) = K10; MU(8 * j + 4) = K11; MU(8 * j + 5) = K12;
        MU(8 * j + 6) = k0;  MU(8 * j + 7) = k1;
        p0 = K00; p1 = K01; p2 = K02; p3 = K10; p4 = K11; p5 = K12; p6 = k0; p7 = k1;
        dec += 0.5 * (qu0 * k0 + qu1 * k1);
        s00 = Q00 - (U00 * K00 + U10 * K10);
        s01 = Q01 - (U00 * K01 + U10 * K11);
        s02 = Q02 - (U00 * K02 + U10 * K12);
        s11 = Q11 - (U01 * K01 + U11 * K11);
        s12 = Q12 - (U01 * K02 + U11 * K12);
        s22 = Q22 - (U02 * K02 + U12 * K12);
        v0 = qz0 - (U00 * k0 + U10 * k1);
        v1 = qz1 - (U01 * k0 + U11 * k1);
        v2 = qz2 - (U02 * k0 + U12 * k1);
    }
    gg_store(m, 0, p0, p1, p2, p3, p4, p5, p6, p7);
    return dec;
}

// gains in U -> the OPEN-LOOP Newton direction of the linear model (K = 0, kf = -d).  The stage constants come from
// the global strip four stages at a time (one L2 round trip per four stages instead of one per stage).
template <int SS>
B2C_HD void slot_open_loop(const MpcArgs& a, const SlotMem& m) {
    const int N = a.N;
    constexpr int C = 4;
    double dx = 0.0, dy = 0.0, dth = 0.0;
    for (int j0 = 0; j0 < N; j0 += C) {
        double ck[C], sk[C], uo0[C], uo1[C];
#pragma unroll
        for (int c = 0; c < C; ++c) g_load4(m, (j0 + c < N) ? (j0 + c) : (N - 1), ck[c], sk[c], uo0[c], uo1[c]);
#pragma unroll
        for (int c = 0; c < C; ++c) {
            const int j = j0 + c;
            if (j < N) {
                const double d0 = -MU(8 * j + 6) - (MU(8 * j + 0) * dx + MU(8 * j + 1) * dy + MU(8 * j + 2) * dth);
                const double d1 = -MU(8 * j + 7) - (MU(8 * j + 3) * dx + MU(8 * j + 4) * dy + MU(8 * j + 5) * dth);
                const double vr = b2c_guard_v(uo0[c], B2C_GUARD_MPC);
                const double ndx = dx + (-vr * sk[c]) * dth + ck[c] * d0;
                const double ndy = dy + (vr * ck[c]) * dth + sk[c] * d0;
                dth = dth + a.dt * d1; dx = ndx; dy = ndy;
#pragma unroll
                for (int e = 0; e < 6; ++e) MU(8 * j + e) = 0.0;
                MU(8 * j + 6) = -d0; MU(8 * j + 7) = -d1;
            }
        }
    }
}

// ----- undo a rejected trial -------------------------------------------------------------------------
// W <- parked previous inputs; U <- gains from the mirror (independent loads, unrolled so they overlap).
// From the MPC_LS_CLOSED-th rejection of an iteration on, the gains are then rewritten as the OPEN-LOOP
// Newton direction of the linear model (d_j = -kf_j - K_j dz_j, dz_{j+1} = A dz_j + B d_j) so that the
// next trial is Bertsekas' projection arc P[w + alpha d], which descends for small alpha because frozen
// inputs are decoupled in Quu.
template <int SS, int MT>
B2C_HD void slot_restore(const MpcArgs& a, const SlotMem& m, const bool open_loop) {
    const int N = a.N;
#pragma unroll 4
    for (int j = 0; j < N; ++j) {
        MW(2 * j) = MU(8 * j + 6); MW(2 * j + 1) = MU(8 * j + 7);
        const double2* p = reinterpret_cast<const double2*>(m.gg + (int64_t)j * 8 * m.gs);
        const double2 q0 = p[0], q1 = p[1], q2 = p[2], q3 = p[3];
        MU(8 * j + 0) = q0.x; MU(8 * j + 1) = q0.y; MU(8 * j + 2) = q1.x;
        MU(8 * j + 3) = q1.y; MU(8 * j + 4) = q2.x; MU(8 * j + 5) = q2.y;
        MU(8 * j + 6) = q3.x; MU(8 * j + 7) = q3.y;
    }
    if (open_loop) slot_open_loop<SS>(a, m);
}

// ----- outputs (:484-522 on success, _get_fallback_solution :316-343 otherwise) -----------------------
template <int SS, int MT>
B2C_HD void slot_finish(const MpcArgs& a, const SlotMem& m, SlotState& s) {
    const int N = a.N;
    const int64_t B = a.B, i = s.inst;
    const double ur00 = s.ur00, ur01 = s.ur01;
    double u00, u01, Jout, mv = s.max_viol;
    const int sc = a.step_count ? a.step_count[i] : a.ramp_up_steps;
    if (s.st == B2C_OK) {
        u00 = ur00 + MW(0); u01 = ur01 + MW(1);
        Jout = s.J;
        if (a.useq || a.xpred) {                  // add the (wrapped) references back, :497-498
            double zx = s.z0x, zy = s.z0y, zt = s.z0t;
            for (int k = 0; k <= N; ++k) {
                if (a.xpred) {
                    a.xpred[(int64_t)(3 * k) * B + i]     = zx + B2C_LDG(a.xref + (int64_t)(3 * k) * B + i);
                    a.xpred[(int64_t)(3 * k + 1) * B + i] = zy + B2C_LDG(a.xref + (int64_t)(3 * k + 1) * B + i);
                    a.xpred[(int64_t)(3 * k + 2) * B + i] = zt + B2C_LDG(a.xref + (int64_t)(3 * k + 2) * B + i);
                }
                if (k == N) break;
                const double wk0 = MW(2 * k), wk1 = MW(2 * k + 1);
                double ck, sk, uo0, uo1;
                g_load4(m, k, ck, sk, uo0, uo1);
                if (a.useq) {
                    a.useq[(int64_t)(2 * k) * B + i] = wk0 + uo0;
                    a.useq[(int64_t)(2 * k + 1) * B + i] = wk1 + uo1;
                }
                const double vr = b2c_guard_v(uo0, B2C_GUARD_MPC);
                const double nzx = zx + (-vr * sk) * zt + ck * wk0, nzy = zy + (vr * ck) * zt + sk * wk0;
                zt = zt + a.dt * wk1; zx = nzx; zy = nzy;
            }
        }
        if (sc < a.ramp_up_steps) {               // cold-start ramp, :500-507
            const double lim = a.w_max * ((double)(sc + 1) / (double)a.ramp_up_steps);
            u01 = b2c_clip(u01, -lim, lim);
            if (a.useq) a.useq[B + i] = u01;
        }
        if (a.step_count) a.step_count[i] = sc + 1;
    } else {
        const double e0 = s.x0x - s.r0x, e2 = b2c_wrap(s.x0t - s.r0t);
        u00 = b2c_clip(ur00 + (-1.0 * e0), -a.v_max, a.v_max);
        u01 = b2c_clip(ur01 + (-0.5 * e2), -a.w_max, a.w_max);
        Jout = (double)INFINITY;
        mv = 0.0;
        if (a.useq || a.xpred) {
            const double x0y = B2C_LDG(a.x0 + B + i);
            for (int k = 0; k <= N; ++k) {
                if (a.xpred) {
                    a.xpred[(int64_t)(3 * k) * B + i] = s.x0x;
                    a.xpred[(int64_t)(3 * k + 1) * B + i] = x0y;
                    a.xpred[(int64_t)(3 * k + 2) * B + i] = s.x0t;
                }
                if (k < N && a.useq) {
                    a.useq[(int64_t)(2 * k) * B + i] = u00;
                    a.useq[(int64_t)(2 * k + 1) * B + i] = u01;
                }
            }
        }
    }
    a.u0[i] = u00; a.u0[B + i] = u01;
    a.cost[i] = Jout;
    if (a.status) a.status[i] = s.st;
    if (a.iters) a.iters[i] = s.it;
    if (a.slack_used) a.slack_used[i] = (mv > 1e-6) ? 1 : 0;   // :485
    s.mode = SLOT_IDLE;
}

// ----- the per-slot state machine: one call = the lane's share of one phase ---------------------------
template <int SS, int MT>
B2C_HD void slot_phase_backward(const MpcArgs& a, const SlotMem& m, SlotState& s) {
    s.dec = slot_backward<SS, MT>(a, m, s);
    if (!(s.dec < 1e300)) { s.st = B2C_NONFINITE; slot_finish<SS, MT>(a, m, s); return; }
    if (s.dec <= a.tol * fmax(1.0, fabs(s.J))) { s.st = B2C_OK; slot_finish<SS, MT>(a, m, s); return; }
    if (s.it >= a.max_iter) { s.st = B2C_MAXITER; slot_finish<SS, MT>(a, m, s); return; }
    s.alpha = (s.it == 0) ? 1.0 : fmin(1.0, MPC_LS_GROW * s.alpha);     // first trial: see mpc_args.cuh
    ++s.it; s.ls = 0; s.mode = SLOT_FWD_GAIN;
}

// what follows a forward sweep.  `krej`: trial steps rejected together when this one is (1 on the host and for a
// lane working alone; 1 + helpers when idle lanes of the warp evaluated the next shorter steps in the same
// round -- the state reached is the one krej successive rejections reach: alpha halves krej times, the
// restore is idempotent).
template <int SS, int MT>
B2C_HD void slot_forward_post(const MpcArgs& a, const SlotMem& m, SlotState& s, const bool gain, const double Jn,
                              const double mv, const double lx, const double ly, const double lt, const int krej) {
    if (!gain) {                                  // first evaluation of a fresh instance
        s.J = Jn; s.max_viol = mv; s.lamx = lx; s.lamy = ly; s.lamt = lt;
        if (!(a.v_max >= 0.0) || !(a.w_max >= 0.0)) { s.st = B2C_INFEASIBLE; slot_finish<SS, MT>(a, m, s); return; }
        if (!(fabs(Jn) < 1e300)) { s.st = B2C_NONFINITE; slot_finish<SS, MT>(a, m, s); return; }
        s.mode = SLOT_BWD;
        return;
    }
    if (Jn < s.J) {                               // accept
        s.J = Jn; s.max_viol = mv; s.lamx = lx; s.lamy = ly; s.lamt = lt; s.mode = SLOT_BWD;
        return;
    }
    for (int r = 0; r < krej; ++r) s.alpha *= 0.5;    // reject: undo, shorten (the arc continues from this alpha)
    s.ls += krej; s.rej = 1;
    slot_restore<SS, MT>(a, m, s.ls >= MPC_LS_CLOSED);
    if (s.ls >= MPC_LS_MAX) {                     // rounding floor or stall
        s.st = (s.dec <= 1e-9 * fmax(1.0, fabs(s.J))) ? B2C_OK : B2C_MAXITER;
        slot_finish<SS, MT>(a, m, s);
    }
}

template <int SS, int MT>
B2C_HD void slot_phase_forward(const MpcArgs& a, const SlotMem& m, SlotState& s) {
    double mv, lx, ly, lt;
    const bool gain = (s.mode == SLOT_FWD_GAIN);
    const double Jn = slot_forward<SS, MT>(a, m, s, gain, mv, lx, ly, lt);
    slot_forward_post<SS, MT>(a, m, s, gain, Jn, mv, lx, ly, lt, 1);
}

// ----- difficulty classes (work ordering only; never changes a result) ------------------------------
// The Newton iteration count of an instance is predicted well by how deep its REFERENCE path runs into
// an inflated obstacle: margin = min_{k<N, m} (|p_ref,k - p_obs,m| - (d_safe + r_m)).  On config 4:
// margin < -0.1 m -> 4.7 iterations on average (max 13), [-0.1, 0) -> 2.9, [0, 0.1) -> 1.7, >= 0.1 -> 1.26
// (max 5).  The persistent kernel claims the hard classes first (longest-processing-time-first), which
// removes the tail a first-come queue leaves when the last instances claimed happen to be slow ones, and
// keeps warps homogeneous (lanes of a warp finish, and refill, together).
#define MPC_NCLASS 4
// Stage loop outside, obstacles in registers (MT of them at a time): every reference position is read once, coalesced
// across the warp, and the MT distance chains are independent -- 21 us for 65,536 instances instead of 34 with the
// obstacle loop outside.  (Four lanes per instance with a shuffle-min: 31 us, and the claim order then groups only 8
// consecutive instances per warp, which costs the refill its coalescing -- measured 0.473 ms per batch against 0.450.)
#ifndef B2C_HOST_EMUL
// min over stages [k0, k1) of the squared distance from the reference position to obstacle m0 + e, e < MT
template <int MT>
__device__ __forceinline__ void mpc_ref_d2min(const MpcArgs& a, int64_t i, int m0, int k0, int k1, double (&d2min)[MT]) {
    const int64_t B = a.B;
    double ox[MT], oy[MT];
#pragma unroll
    for (int e = 0; e < MT; ++e) {
        const int m = (m0 + e < a.M) ? (m0 + e) : (a.M - 1);
        ox[e] = __ldg(a.obs + (int64_t)(3 * m) * B + i); oy[e] = __ldg(a.obs + (int64_t)(3 * m + 1) * B + i);
        d2min[e] = 1e300;
    }
#pragma unroll 5
    for (int k = k0; k < k1; ++k) {
        const double px = __ldg(a.xref + (int64_t)(3 * k) * B + i), py = __ldg(a.xref + (int64_t)(3 * k + 1) * B + i);
#pragma unroll
        for (int e = 0; e < MT; ++e) {
            const double dx = px - ox[e], dy = py - oy[e];
            d2min[e] = fmin(d2min[e], dx * dx + dy * dy);
        }
    }
}
#endif

#ifndef B2C_HOST_EMUL
// perm[c*B + pos] = instance ids of class c in arrival order.  Counters are aggregated per CTA of 256 instances (one
// atomicAdd per class and CTA: 2,048 warps adding to the same four words serialised at L2 and were most of the
// kernel's 21 us), so a CTA's instances stay contiguous and in index order inside each class.
#define MPC_CLASSIFY_THREADS 512
#define MPC_CLASSIFY_INST 256                         // instances per CTA: two threads each (first / second half of the horizon)
__global__ void __launch_bounds__(MPC_CLASSIFY_THREADS)
k_mpc_classify(const MpcArgs a) {
    constexpr int MT = 5;
    __shared__ double half2[MT][MPC_CLASSIFY_INST];                        // second half's minima
    __shared__ unsigned wcount[MPC_CLASSIFY_INST / 32][MPC_NCLASS];        // per warp, then exclusive prefix over warps
    __shared__ unsigned long long cbase[MPC_NCLASS];
    const int li = threadIdx.x & (MPC_CLASSIFY_INST - 1), half = threadIdx.x / MPC_CLASSIFY_INST;
    const int64_t i = (int64_t)blockIdx.x * MPC_CLASSIFY_INST + li, B = a.B;
    const bool valid = i < B;
    const int lane = threadIdx.x & 31, warp = li >> 5;
    const int kh = a.N / 2;
    int cls = MPC_NCLASS - 1;
    for (int m0 = 0; m0 < a.M; m0 += MT) {             // uniform trip count: barriers inside
        double d2min[MT];
#pragma unroll
        for (int e = 0; e < MT; ++e) d2min[e] = 1e300;
        if (valid) mpc_ref_d2min<MT>(a, i, m0, half ? kh : 0, half ? a.N : kh, d2min);
        if (half) {
#pragma unroll
            for (int e = 0; e < MT; ++e) half2[e][li] = d2min[e];
        }
        __syncthreads();
        if (!half && valid) {
#pragma unroll
            for (int e = 0; e < MT; ++e) {
                const double d2 = fmin(d2min[e], half2[e][li]);
                const int m = (m0 + e < a.M) ? (m0 + e) : (a.M - 1);
                const double c = a.d_safe + __ldg(a.obs + (int64_t)(3 * m + 2) * B + i);
                const double cm = fmax(c - 0.1, 0.0), cp = c + 0.1;
                const double t0 = cm * cm, t1 = c * fabs(c), t2 = cp * fabs(cp);
                const int cm_ = (d2 < t0) ? 0 : (d2 < t1) ? 1 : (d2 < t2) ? 2 : 3;
                cls = (cm_ < cls) ? cm_ : cls;
            }
        }
        __syncthreads();
    }
    if (half || !valid) cls = -1;
    unsigned mymask = 0;
#pragma unroll
    for (int c = 0; c < MPC_NCLASS; ++c) {
        const unsigned mask = __ballot_sync(0xffffffffu, cls == c);
        if (cls == c) mymask = mask;
        if (lane == 0 && !half) wcount[warp][c] = __popc(mask);
    }
    __syncthreads();
    if (threadIdx.x < MPC_NCLASS) {
        const int c = threadIdx.x;
        unsigned tot = 0;
        for (int w = 0; w < MPC_CLASSIFY_INST / 32; ++w) { const unsigned n = wcount[w][c]; wcount[w][c] = tot; tot += n; }
        cbase[c] = tot ? atomicAdd(a.class_count + c, (unsigned long long)tot) : 0ull;
    }
    __syncthreads();
    if (cls >= 0)
        a.perm[(int64_t)cls * B + (int64_t)cbase[cls] + wcount[warp][cls] + __popc(mymask & ((1u << lane) - 1u))] = (uint32_t)i;
}
#endif

// ----- kernel ---------------------------------------------------------------------------------------
#ifndef B2C_HOST_EMUL
#ifdef B2C_SLOT_TRACE                                 // tools/ only (make trace): per-warp phase timing
__device__ unsigned long long g_slot_trace[16 * 4096];
__device__ __forceinline__ unsigned long long trace_ns() {
    unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t;
}
#define TRACE(...) __VA_ARGS__
#else
#define TRACE(...)
#endif
// Helper lanes (the tail of a batch).  The kernel's makespan at BASELINE's batch size is the latency of its
// hardest instances: 13 Newton iterations plus as many rejected trial steps, one forward sweep each.  An instance
// that has had a trial rejected (s.rej) therefore keeps MPC_SPEC_HELPERS (or all, see MPC_SPEC_FEW) idle lanes of its warp from
// refilling while the queue is short (remaining <= MPC_SPEC_TAIL x resident slots), and those lanes evaluate the
// NEXT trial steps of the same iteration (alpha/2, alpha/4, ...) in the same round, on copies of the owner's strips.
// The first trial of the sequence that decreases the cost is taken -- the sequence and every sweep are those of
// the sequential line search, so results do not depend on whether, or how many, helpers were there.  If the winner
// is a helper, the instance simply continues on that lane (state by shuffle, the two lanes swap global strips).
#ifndef MPC_SPEC_HELPERS
#define MPC_SPEC_HELPERS 8
#endif
#ifndef MPC_SPEC_TAIL
#define MPC_SPEC_TAIL 4
#endif
// a warp with at most MPC_SPEC_FEW such instances keeps ALL its idle lanes back (it stops refilling: its rounds lose the
// refill phase and every trial step of an iteration is evaluated at once); with more of them, MPC_SPEC_HELPERS lanes each.
// Measured on config 4 (ms per 65,536 batch): 3 helpers flat 0.441, 8 flat 0.426, 31 flat 0.405 -- but a flat 31 makes the
// closed-loop rollouts (N = 6, many instances with rejected trials per warp) bimodal, 0.32 s or 0.95 s per config-5 shard,
// because whole warps stall behind their requesters; FEW = 2 with 8 helpers: 0.410 and the shard stays at 0.315 s.
#ifndef MPC_SPEC_FEW
#define MPC_SPEC_FEW 2
#endif
template <int SS, int MT>
__global__ void __launch_bounds__(SS, 1)
k_mpc_slots(const MpcArgs a) {
    extern __shared__ double smem[];
    constexpr int tpb = SS;
    const int lane = threadIdx.x & 31;
    const unsigned lt_mask = (1u << lane) - 1u;
    int64_t strip = (int64_t)blockIdx.x * tpb + threadIdx.x;  // global strip this lane's instance uses (moves with it)
    double* const gg_base = a.scratch + (int64_t)6 * a.N * a.nslots;
    SlotMem m;
    m.w = smem + threadIdx.x;
    m.u = m.w + (size_t)2 * a.N * tpb;
    m.ob = m.u + (size_t)8 * a.N * tpb;
    // ring memory: 12 x 32 doubles per WARP (the two layouts below overlap inside a warp's block only -- warps are not
    // synchronised with each other, a warp's own phases are)
    double* const wring = smem + (size_t)slot_fields(a) * tpb + (size_t)(threadIdx.x >> 5) * 384;
    m.ring = wring + (size_t)lane * 6;                        // forward: 2 stages x [32 lanes][6]
    m.ring4 = wring + (size_t)lane * 4;                       // backward: 3 stages x [32 lanes][4]
    m.gs = a.nslots;
    m.g = a.scratch + strip * 6;
    m.gg = gg_base + strip * 8;
    SlotState s;
    s.mode = SLOT_IDLE; s.inst = -1; s.it = 0; s.ls = 0; s.st = 0; s.rej = 0;
    s.alpha = s.J = s.dec = s.z0x = s.z0y = s.z0t = s.lamx = s.lamy = s.lamt = s.max_viol = 0.0;
    s.x0x = s.x0t = s.r0x = s.r0t = s.ur00 = s.ur01 = 0.0;
    bool empty = false;
    const unsigned FULL = 0xffffffffu;
    unsigned long long cum0 = 0, cum1 = 0, cum2 = 0;          // class boundaries in queue order
    if (a.perm) { cum0 = a.class_count[0]; cum1 = cum0 + a.class_count[1]; cum2 = cum1 + a.class_count[2]; }
    int64_t seen = 0;                                         // queue position at this warp's last claim
    const int nfields = slot_fields(a);
    TRACE(unsigned long long tr_t0 = trace_ns(); long long tr_b = 0, tr_r = 0, tr_f = 0, tr_c, tr_rounds = 0, tr_lanes = 0;
          long long tr_x0 = 0, tr_x1 = 0, tr_x2 = 0, tr_x3 = 0, tr_x4 = 0, tr_x5 = 0;)
    for (;;) {
        TRACE(tr_c = clock64(); ++tr_rounds; tr_lanes += __popc(__ballot_sync(FULL, s.mode != SLOT_IDLE));)
        if (s.mode == SLOT_BWD) slot_phase_backward<SS, MT>(a, m, s);
        __syncwarp();
        TRACE(tr_b += clock64() - tr_c; tr_c = clock64();)
        // lanes kept back as helpers: MPC_SPEC_HELPERS per instance with a rejected trial behind it, near the tail only
        const unsigned idle0 = __ballot_sync(FULL, s.mode == SLOT_IDLE);
        const unsigned req_m = __ballot_sync(FULL, s.mode == SLOT_FWD_GAIN && s.rej);
        int keep = 0;
        if (req_m && (a.B - seen) <= (int64_t)MPC_SPEC_TAIL * a.nslots) {
            const int nreq = __popc(req_m);
            keep = ((nreq <= MPC_SPEC_FEW) ? 31 : MPC_SPEC_HELPERS) * nreq;
        }
        const bool want = (s.mode == SLOT_IDLE) && !empty && (__popc(idle0 & lt_mask) < __popc(idle0) - keep);
        const unsigned need = __ballot_sync(FULL, want);
        if (need) {
            const int leader = __ffs(need) - 1;
            unsigned long long base = 0;
            if (lane == leader) base = atomicAdd(a.queue, (unsigned long long)__popc(need));
            base = __shfl_sync(FULL, base, leader);
            seen = (int64_t)base + __popc(need);
            int64_t id = 0;
            bool id_ok = false;
            if (want) {
                id = (int64_t)base + __popc(need & lt_mask);
                if (id < a.B) {
                    id_ok = true;
                    if (a.perm) {                             // queue position -> (class, offset) -> instance
                        const unsigned long long q = (unsigned long long)id;
                        const int c = (q >= cum0) + (q >= cum1) + (q >= cum2);
                        const unsigned long long lo = (c == 0) ? 0ull : (c == 1) ? cum0 : (c == 2) ? cum1 : cum2;
                        id = (int64_t)a.perm[(int64_t)c * a.B + (int64_t)(q - lo)];
                    }
                }
            }
            const bool fresh = want && id_ok;
            TRACE(tr_x0 += clock64() - tr_c; long long tr_c2 = clock64();)
            // the refilling lanes' setup, shared by the whole warp: task t = (stage k, refilling lane n),
            // n fastest so that neighbouring tasks read neighbouring instances of the same input row.
            // (k, n) advance by (32 / cnt, 32 % cnt) per pass; lane r holds the r-th refilling lane's number.
            const unsigned vmask = __ballot_sync(FULL, fresh);
            const int cnt = __popc(vmask), T = cnt * a.N;
            if (cnt) {
                const int nth = __fns(vmask, 0, (lane < cnt ? lane : 0) + 1);
                const int dk = 32 / cnt, dn = 32 - dk * cnt;
                // the fresh lanes' own scalars and obstacle rows go first: their latency hides behind the row loop
                double ox0x = 0.0, ox0y = 0.0, ox0t = 0.0, or0x = 0.0, or0y = 0.0, or0t = 0.0;
                if (fresh) {
                    const int64_t B = a.B;
                    ox0x = __ldg(a.x0 + id); ox0y = __ldg(a.x0 + B + id); ox0t = __ldg(a.x0 + 2 * B + id);
                    or0x = __ldg(a.xref + id); or0y = __ldg(a.xref + B + id); or0t = __ldg(a.xref + 2 * B + id);
                    const int M3 = 3 * ((MT >= 0) ? MT : a.M);
                    const unsigned dst = (unsigned)__cvta_generic_to_shared(&MO(0));
                    for (int e = 0; e < M3; ++e)
                        asm volatile("cp.async.ca.shared.global [%0], [%1], 8;"
                                     :: "r"(dst + (unsigned)(e * SS * 8)), "l"(a.obs + (int64_t)e * B + id) : "memory");
                }
                int k = lane / cnt, n = lane - k * cnt;
                for (int t = lane; t - lane < T; t += 32) {
                    const int tl = __shfl_sync(FULL, nth, n);
                    const int64_t tid = __shfl_sync(FULL, id, tl);
                    if (t < T) {
                        SlotMem tm = m;
                        tm.w += tl - lane; tm.u += tl - lane;
                        setup_load_stage_async<SS>(a, tm, tid, k);
                    }
                    k += dk; n += dn;
                    if (n >= cnt) { n -= cnt; ++k; }
                }
                TRACE(tr_x4 += clock64() - tr_c2;)
                if (fresh) {                                      // the scalars loaded before the row loop
                    s.inst = id; s.it = 0; s.ls = 0; s.rej = 0; s.st = B2C_MAXITER; s.alpha = 0.0; s.dec = 0.0; s.max_viol = 0.0;
                    s.x0x = ox0x; s.x0t = ox0t; s.r0x = or0x; s.r0t = or0t;
                    s.z0x = ox0x - or0x; s.z0y = ox0y - or0y; s.z0t = b2c_wrap(ox0t - or0t);       // :392-401, :421
                    s.mode = SLOT_FWD_PLAIN;
                }
                TRACE(__syncwarp(); tr_x5 += clock64() - tr_c2;)
                asm volatile("cp.async.wait_all;" ::: "memory");
                __syncwarp();
                TRACE(tr_x1 += clock64() - tr_c2; tr_c2 = clock64();)
                // np.unwrap: each fresh lane looks along its own column for a heading step of pi or more (np.unwrap is
                // the identity otherwise); only then the sequential pass.  The obstacle strip gets its d_safe + r.
                bool jump = false;
                if (fresh) {
                    const int M3 = 3 * ((MT >= 0) ? MT : a.M);
                    for (int e = 2; e < M3; e += 3) MO(e) = a.d_safe + MO(e);
                    double prev = MU(0);
#pragma unroll 4
                    for (int kk = 1; kk < a.N; ++kk) {
                        const double cur = MU(8 * kk);
                        jump |= !(fabs(cur - prev) < 3.141592653589793);
                        prev = cur;
                    }
                }
                const bool any_jump = __any_sync(FULL, jump);
                if (fresh) setup_unwrap<SS>(a, m, s, any_jump);
                __syncwarp();
                TRACE(tr_x2 += clock64() - tr_c2; tr_c2 = clock64();)
                k = lane / cnt; n = lane - k * cnt;
                for (int t = lane; t - lane < T; t += 32) {
                    const int tl = __shfl_sync(FULL, nth, n);
                    const int64_t tstrip = __shfl_sync(FULL, strip, tl);
                    if (t < T) {
                        SlotMem tm = m;
                        tm.w += tl - lane; tm.u += tl - lane; tm.g = a.scratch + tstrip * 6;
                        setup_stage<SS>(a, tm, k);
                    }
                    k += dk; n += dn;
                    if (n >= cnt) { n -= cnt; ++k; }
                }
                TRACE(__syncwarp(); tr_x3 += clock64() - tr_c2;)
            }
            if ((int64_t)base + __popc(need) >= a.B) empty = true;
        }
        if (empty && __all_sync(FULL, s.mode == SLOT_IDLE)) break;
        __syncwarp();
        TRACE(tr_r += clock64() - tr_c; tr_c = clock64();)
        // ---- forward phase: every lane with a sweep to run, plus the idle lanes lent to requesting instances
        const bool gain = (s.mode == SLOT_FWD_GAIN);
        const unsigned idle_m = __ballot_sync(FULL, s.mode == SLOT_IDLE);
        const bool team_round = (idle_m != 0u) && (req_m != 0u);          // warp-uniform
        int ol = lane, off = 0, ols = 0;                                  // team owner, trial offset, owner's trial index
        bool helper = false;
        SlotState fs = s;
        SlotMem mf = m;
        double Jref = s.J;
        if (team_round) {
            const int nown = __popc(req_m);
            const int ir = __popc(idle_m & lt_mask);
            const int cand = __fns(req_m, 0, (ir % nown) + 1);
            const int coff = 1 + ir / nown;
            ols = __shfl_sync(FULL, s.ls, cand);
            helper = (s.mode == SLOT_IDLE) && (ols + coff <= MPC_LS_MAX - 1);
            if (helper) { ol = cand; off = coff; }
            const double oz0x = __shfl_sync(FULL, s.z0x, ol), oz0y = __shfl_sync(FULL, s.z0y, ol);
            const double oz0t = __shfl_sync(FULL, s.z0t, ol), oal = __shfl_sync(FULL, s.alpha, ol);
            const double oJ = __shfl_sync(FULL, s.J, ol);
            const int64_t ostrip = __shfl_sync(FULL, strip, ol);
            if (helper) {
                fs.z0x = oz0x; fs.z0y = oz0y; fs.z0t = oz0t; Jref = oJ;
                double al = oal;
                for (int r = 0; r < off; ++r) al *= 0.5;
                fs.alpha = al;
                mf.g = a.scratch + ostrip * 6;
                const int d = ol - lane;                                  // the owner's strips (W, U, OB) -> mine
                int f = 0;                                                // eight loads in flight per store batch
                for (; f + 8 <= nfields; f += 8) {
                    double t8[8];
#pragma unroll
                    for (int e = 0; e < 8; ++e) t8[e] = m.w[(size_t)(f + e) * SS + d];
#pragma unroll
                    for (int e = 0; e < 8; ++e) m.w[(size_t)(f + e) * SS] = t8[e];
                }
                for (; f < nfields; ++f) m.w[(size_t)f * SS] = m.w[(size_t)f * SS + d];
                if (ols < MPC_LS_CLOSED && ols + off >= MPC_LS_CLOSED) slot_open_loop<SS>(a, mf);
            }
            __syncwarp();
        }
        const bool run = gain || (s.mode == SLOT_FWD_PLAIN) || helper;
        const bool fgain = gain || helper;
        double Jn = 0.0, mv = 0.0, lx = 0.0, ly = 0.0, lt = 0.0;
        if (run) Jn = slot_forward<SS, MT>(a, mf, fs, fgain, mv, lx, ly, lt);
        __syncwarp();
        int krej = 1;
        bool handled = false;                                             // my instance moved to a helper's lane
        if (team_round) {
            const unsigned okb = __ballot_sync(FULL, run && fgain && (Jn < Jref));
            const unsigned team = __match_any_sync(FULL, ol);
            const unsigned tok = team & okb;
            int win = -1;                                                 // lane whose trial is taken
            if (tok & (1u << ol)) win = ol;
            else if (tok) win = __ffs(tok) - 1;                           // helpers: offsets ascend with the lane number
            krej = __popc(team);
            const bool moved = (win >= 0 && win != ol);
            if (__any_sync(FULL, moved)) {
                const int src = (moved && lane == win) ? ol : lane;       // new owner reads the old owner's state
                const int par = moved ? ((lane == win) ? ol : ((lane == ol) ? win : lane)) : lane;
                const int64_t pstrip = __shfl_sync(FULL, strip, par);
                SlotState t;
                t.inst = __shfl_sync(FULL, s.inst, src);
                t.it = __shfl_sync(FULL, s.it, src); t.st = __shfl_sync(FULL, s.st, src);
                t.dec = __shfl_sync(FULL, s.dec, src);
                t.z0x = __shfl_sync(FULL, s.z0x, src); t.z0y = __shfl_sync(FULL, s.z0y, src); t.z0t = __shfl_sync(FULL, s.z0t, src);
                t.x0x = __shfl_sync(FULL, s.x0x, src); t.x0t = __shfl_sync(FULL, s.x0t, src);
                t.r0x = __shfl_sync(FULL, s.r0x, src); t.r0t = __shfl_sync(FULL, s.r0t, src);
                t.ur00 = __shfl_sync(FULL, s.ur00, src); t.ur01 = __shfl_sync(FULL, s.ur01, src);
                if (moved && lane == win) {                               // I take the instance over, trial accepted
                    s.inst = t.inst; s.it = t.it; s.st = t.st; s.dec = t.dec;
                    s.z0x = t.z0x; s.z0y = t.z0y; s.z0t = t.z0t; s.x0x = t.x0x; s.x0t = t.x0t; s.r0x = t.r0x; s.r0t = t.r0t;
                    s.ur00 = t.ur00; s.ur01 = t.ur01;
                    s.alpha = fs.alpha; s.ls = ols + off; s.rej = 1;
                    s.J = Jn; s.max_viol = mv; s.lamx = lx; s.lamy = ly; s.lamt = lt;
                    s.mode = SLOT_BWD;
                    strip = pstrip;
                } else if (moved && lane == ol) {                         // my instance went to lane `win`
                    s.mode = SLOT_IDLE; s.inst = -1; s.rej = 0;
                    strip = pstrip;
                    handled = true;
                }
                m.g = a.scratch + strip * 6;
                m.gg = gg_base + strip * 8;
            }
        }
        if (!handled && !helper && run) slot_forward_post<SS, MT>(a, m, s, gain, Jn, mv, lx, ly, lt, krej);
        __syncwarp();
        TRACE(tr_f += clock64() - tr_c;)
    }
    TRACE(if (lane == 0) {
        const int w = (blockIdx.x * SS + threadIdx.x) >> 5;
        if (w < 4096) {
            unsigned long long* t = g_slot_trace + 16 * w;
            t[0] = tr_t0; t[1] = trace_ns(); t[2] = tr_rounds; t[3] = tr_b; t[4] = tr_r; t[5] = tr_f; t[6] = tr_lanes; t[7] = 1;
            t[8] = tr_x0; t[9] = tr_x1; t[10] = tr_x2; t[11] = tr_x3; t[12] = tr_x4; t[13] = tr_x5;
        }
    })
}
#endif

}  // namespace


bool mpc_slots_eligible(const MpcArgs& a, int* tpb_out) {
    if (a.kind != B2C_MPC_LTV || a.bs != 1 || !a.soft) return false;
    const size_t cap = 227 * 1024;
    for (int tpb = 128; tpb >= 32; tpb >>= 1) {
        if ((size_t)(slot_fields(a) + 12) * tpb * sizeof(double) <= cap) { *tpb_out = tpb; return true; }
    }
    return false;
}

#ifndef B2C_HOST_EMUL
template <int SS, int MT>
static int slots_launch_t(b2c_ctx* ctx, const MpcArgs& a, int grid, size_t smem) {
    B2C_CUDA(cudaFuncSetAttribute(k_mpc_slots<SS, MT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    k_mpc_slots<SS, MT><<<grid, SS, smem, ctx->stream>>>(a);
    B2C_CHECK_LAUNCH(ctx);
    return 0;
}

int mpc_slots_launch(b2c_ctx* ctx, MpcArgs& a, int tpb) {
    const size_t smem = (size_t)(slot_fields(a) + 12) * tpb * sizeof(double);
    int grid = (int)((a.B + tpb - 1) / tpb);
    // short horizons leave room for a second resident CTA per SM (two warps per scheduler hide the dependent-issue
    // latency of the sweeps); registers allow two CTAs of 128 threads, shared memory decides
    const int per_sm = (tpb == 128 && 2 * (smem + 1024) <= (size_t)227 * 1024) ? 2 : 1;
    if (grid > per_sm * ctx->sm_count) grid = per_sm * ctx->sm_count;
    a.nslots = (int64_t)grid * tpb;
    // global strips: G[6N] + GG[8N] doubles per slot, then the queue counter
    const size_t strip = (size_t)(14 * a.N) * (size_t)a.nslots * sizeof(double);
    // more than one instance per slot: claim in difficulty order (see mpc_difficulty_class)
    const bool ordered = a.M > 0 && a.B > a.nslots && a.B < ((int64_t)1 << 32);
    const size_t perm_bytes = ordered ? (size_t)MPC_NCLASS * (size_t)a.B * sizeof(uint32_t) : 0;
    int rc = b2c_reserve(&ctx->scratch, &ctx->scratch_bytes, strip + 256 + perm_bytes, false);
    if (rc) return rc;
    a.scratch = (double*)ctx->scratch;
    a.queue = (unsigned long long*)((char*)ctx->scratch + strip);
    a.class_count = a.queue + 1;
    a.perm = nullptr;
    B2C_CUDA(cudaMemsetAsync(a.queue, 0, (1 + MPC_NCLASS) * sizeof(unsigned long long), ctx->stream));
    if (ordered) {
        a.perm = (uint32_t*)((char*)ctx->scratch + strip + 256);
        k_mpc_classify<<<b2c_blocks(a.B, MPC_CLASSIFY_INST), MPC_CLASSIFY_THREADS, 0, ctx->stream>>>(a);
        B2C_CHECK_LAUNCH(ctx);
    }
    // obstacle rows unrolled at compile time for the two shapes the reference runs (M = 3, 5)
    if (tpb == 128) {
        if (a.M == 5) return slots_launch_t<128, 5>(ctx, a, grid, smem);
        if (a.M == 3) return slots_launch_t<128, 3>(ctx, a, grid, smem);
        if (a.M == 0) return slots_launch_t<128, 0>(ctx, a, grid, smem);
        return slots_launch_t<128, -1>(ctx, a, grid, smem);
    }
    if (tpb == 64) return slots_launch_t<64, -1>(ctx, a, grid, smem);
    return slots_launch_t<32, -1>(ctx, a, grid, smem);
}
#ifdef B2C_SLOT_TRACE
extern "C" int b2c_debug_slot_trace(unsigned long long* out, int nwarps) {
    B2C_CUDA(cudaMemcpyFromSymbol(out, g_slot_trace, (size_t)nwarps * 16 * sizeof(unsigned long long)));
    return 0;
}
#endif
#else
int mpc_slots_launch(b2c_ctx*, MpcArgs&, int) { return -1; }

#include <vector>
template <int MT>
static void emul_one(const MpcArgs& a, const SlotMem& m, int64_t i) {
    constexpr int SS = 1;
    SlotState s;
    slot_setup<SS, MT>(a, m, s, i);
    while (s.mode != SLOT_IDLE) {
        if (s.mode == SLOT_BWD) slot_phase_backward<SS, MT>(a, m, s);
        if (s.mode == SLOT_FWD_GAIN || s.mode == SLOT_FWD_PLAIN) slot_phase_forward<SS, MT>(a, m, s);
    }
}
// Test-only: the slot state machine on the host, one instance at a time (SoA host pointers).
extern "C" int b2c_emul_mpc_slots(int64_t B, const b2c_mpc_params* p,
                                  const double* x0, const double* xref, const double* uref, const double* obs,
                                  int32_t* step_count, double* u0, double* cost, int32_t* status,
                                  int32_t* iters, int32_t* slack_used, double* useq, double* xpred) {
    MpcArgs a;
    mpc_fill_args(a, B, p, x0, xref, uref, obs, step_count, u0, cost, status, iters, slack_used, useq, xpred);
    int tpb = 0;
    if (!mpc_slots_eligible(a, &tpb)) return -5;
    std::vector<double> sm((size_t)slot_fields(a)), gl((size_t)14 * a.N);
    SlotMem m;
    m.ring = nullptr; m.ring4 = nullptr; m.w = sm.data(); m.u = m.w + 2 * a.N; m.ob = m.u + 8 * a.N;
    m.gs = 1; m.g = gl.data(); m.gg = m.g + 6 * a.N;
    for (int64_t i = 0; i < B; ++i) {
        if (a.M == 5) emul_one<5>(a, m, i);
        else if (a.M == 3) emul_one<3>(a, m, i);
        else if (a.M == 0) emul_one<0>(a, m, i);
        else emul_one<-1>(a, m, i);
    }
    return 0;
}
#endif
