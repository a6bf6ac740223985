// K1: batched Jacobian + discretization; K2: batched fp64 DARE (structured doubling) + LQR gain;
// LQR control law.  One thread per instance, everything in registers, SoA coalesced I/O.
//
// Reference behaviour being reproduced (HC = /root/reference/src/hybrid_controller/hybrid_controller):
//   HC/models/linearization.py:62-225, HC/controllers/lqr_controller.py:92-257.
#include "b2c_common.cuh"

// ---------------------------------------------------------------------------------------------
// K1 core: the 4 (euler/jacobian) or 6 (exact) non-trivial entries at one operating point.
// Operation order matches numpy's so the only difference to the reference is sincos vs libm.
struct LinModel {
    double a13, a23;          // A[0][2], A[1][2]
    double b11, b21;          // B[0][0], B[1][0]
    double b12, b22;          // B[0][1], B[1][1]  (non-zero only for EXACT)
    double b32;               // B[2][1]
    double adiag;             // 1 (discrete) or 0 (continuous Jacobian)
};

B2C_HD LinModel b2c_linearize_point(double v, double th, double dt, int method) {
    double s, c;
    sincos(th, &s, &c);
    LinModel m;
    if (method == B2C_LIN_JACOBIAN) {             // linearization.py:62-96
        m.a13 = -v * s; m.a23 = v * c;
        m.b11 = c; m.b21 = s; m.b12 = 0.0; m.b22 = 0.0; m.b32 = 1.0; m.adiag = 0.0;
    } else {                                      // :98-118 == :190-225
        m.a13 = B2C_MUL(B2C_MUL(-v, s), dt);
        m.a23 = B2C_MUL(B2C_MUL(v, c), dt);
        m.b11 = B2C_MUL(c, dt); m.b21 = B2C_MUL(s, dt);
        m.b12 = 0.0; m.b22 = 0.0; m.b32 = dt; m.adiag = 1.0;
        if (method == B2C_LIN_EXACT) {            // :120-158; A^2 = 0 => series terminates:
            // A_d = I + A dt,  B_d = B dt + A B dt^2/2
            double h = 0.5 * dt * dt;
            m.b12 = B2C_MUL(B2C_MUL(-v, s), h);
            m.b22 = B2C_MUL(B2C_MUL(v, c), h);
        }
    }
    return m;
}

__global__ void __launch_bounds__(256)
k_linearize_soa(int64_t B, int method, int guard, double dt,
                const double* __restrict__ v, const double* __restrict__ th,
                double* __restrict__ Ad, double* __restrict__ Bd) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B) return;
    LinModel m = b2c_linearize_point(b2c_guard_v(v[i], guard), th[i], dt, method);
    Ad[0 * B + i] = m.adiag; Ad[1 * B + i] = 0.0;     Ad[2 * B + i] = m.a13;
    Ad[3 * B + i] = 0.0;     Ad[4 * B + i] = m.adiag; Ad[5 * B + i] = m.a23;
    Ad[6 * B + i] = 0.0;     Ad[7 * B + i] = 0.0;     Ad[8 * B + i] = m.adiag;
    Bd[0 * B + i] = m.b11;   Bd[1 * B + i] = m.b12;
    Bd[2 * B + i] = m.b21;   Bd[3 * B + i] = m.b22;
    Bd[4 * B + i] = 0.0;     Bd[5 * B + i] = m.b32;
}

// numpy layout (B,3,3)/(B,3,2): stage a CTA's tile in shared memory, write it out contiguously.
#define LIN_TPB 256
__global__ void __launch_bounds__(LIN_TPB)
k_linearize_packed(int64_t B, int method, int guard, double dt,
                   const double* __restrict__ v, const double* __restrict__ th,
                   double* __restrict__ Ad, double* __restrict__ Bd) {
    __shared__ double sA[LIN_TPB * 9];
    __shared__ double sB[LIN_TPB * 6];
    int64_t base = (int64_t)blockIdx.x * LIN_TPB;
    int t = threadIdx.x;
    int64_t i = base + t;
    if (i < B) {
        LinModel m = b2c_linearize_point(b2c_guard_v(v[i], guard), th[i], dt, method);
        double* a = sA + t * 9; double* b = sB + t * 6;
        a[0] = m.adiag; a[1] = 0.0; a[2] = m.a13;
        a[3] = 0.0; a[4] = m.adiag; a[5] = m.a23;
        a[6] = 0.0; a[7] = 0.0; a[8] = m.adiag;
        b[0] = m.b11; b[1] = m.b12; b[2] = m.b21; b[3] = m.b22; b[4] = 0.0; b[5] = m.b32;
    }
    __syncthreads();
    int64_t nvalid = B - base; if (nvalid > LIN_TPB) nvalid = LIN_TPB;
    double* oa = Ad + base * 9; double* ob = Bd + base * 6;
    for (int j = t; j < nvalid * 9; j += LIN_TPB) oa[j] = sA[j];
    for (int j = t; j < nvalid * 6; j += LIN_TPB) ob[j] = sB[j];
}

int b2c_linearize_dev(b2c_ctx* ctx, int64_t B, int method, int guard, double dt,
                      const double* v, const double* th, double* Ad, double* Bd) {
    if (!ctx) { b2c_set_error("null context"); return -1; }
    if (method < 0 || method > 2) { b2c_set_error("Unknown discretization method"); return -2; }
    if (B <= 0) return 0;
    B2C_DEVICE_GUARD(ctx);
    k_linearize_soa<<<b2c_blocks(B, 256), 256, 0, ctx->stream>>>(B, method, guard, dt, v, th, Ad, Bd);
    B2C_CHECK_LAUNCH(ctx);
    return 0;
}

int b2c_linearize_packed_dev(b2c_ctx* ctx, int64_t B, int method, int guard, double dt,
                             const double* v, const double* th, double* Ad, double* Bd) {
    if (!ctx) { b2c_set_error("null context"); return -1; }
    if (method < 0 || method > 2) { b2c_set_error("Unknown discretization method"); return -2; }
    if (B <= 0) return 0;
    B2C_DEVICE_GUARD(ctx);
    k_linearize_packed<<<b2c_blocks(B, LIN_TPB), LIN_TPB, 0, ctx->stream>>>(B, method, guard, dt, v, th, Ad, Bd);
    B2C_CHECK_LAUNCH(ctx);
    return 0;
}

// ---------------------------------------------------------------------------------------------
// K2: DARE by the structure-preserving doubling algorithm (SDA), 3x3 blocks in registers.
//   A_0 = A_d, G_0 = B R^-1 B', H_0 = Q
//   W = I + G H;  A+ = A W^-1 A;  G+ = G + A W^-1 G A';  H+ = H + A' H W^-1 A;   H -> P
// Quadratic convergence; the closed loop of this plant has |lambda| ~ 0.8-0.99 => 9-17 steps.
// Replaces scipy.linalg.solve_discrete_are (QZ on the symplectic pencil), lqr_controller.py:126.
struct M3 { double m[9]; };

B2C_HD M3 m3_mul(const M3& a, const M3& b) {
    M3 r;
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++)
            r.m[i * 3 + j] = fma(a.m[i * 3 + 2], b.m[6 + j], fma(a.m[i * 3 + 1], b.m[3 + j], a.m[i * 3] * b.m[j]));
    return r;
}
B2C_HD M3 m3_mul_tn(const M3& a, const M3& b) {   // a' b
    M3 r;
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++)
            r.m[i * 3 + j] = fma(a.m[6 + i], b.m[6 + j], fma(a.m[3 + i], b.m[3 + j], a.m[i] * b.m[j]));
    return r;
}
B2C_HD M3 m3_mul_nt(const M3& a, const M3& b) {   // a b'
    M3 r;
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++)
            r.m[i * 3 + j] = fma(a.m[i * 3 + 2], b.m[j * 3 + 2], fma(a.m[i * 3 + 1], b.m[j * 3 + 1], a.m[i * 3] * b.m[j * 3]));
    return r;
}
B2C_HD M3 m3_inv(const M3& a) {
    double c00 = fma(a.m[4], a.m[8], -a.m[5] * a.m[7]);
    double c01 = fma(a.m[5], a.m[6], -a.m[3] * a.m[8]);
    double c02 = fma(a.m[3], a.m[7], -a.m[4] * a.m[6]);
    double det = fma(a.m[0], c00, fma(a.m[1], c01, a.m[2] * c02));
    double id = 1.0 / det;
    M3 r;
    r.m[0] = c00 * id;
    r.m[1] = fma(a.m[2], a.m[7], -a.m[1] * a.m[8]) * id;
    r.m[2] = fma(a.m[1], a.m[5], -a.m[2] * a.m[4]) * id;
    r.m[3] = c01 * id;
    r.m[4] = fma(a.m[0], a.m[8], -a.m[2] * a.m[6]) * id;
    r.m[5] = fma(a.m[2], a.m[3], -a.m[0] * a.m[5]) * id;
    r.m[6] = c02 * id;
    r.m[7] = fma(a.m[1], a.m[6], -a.m[0] * a.m[7]) * id;
    r.m[8] = fma(a.m[0], a.m[4], -a.m[1] * a.m[3]) * id;
    return r;
}
// products whose result is symmetric in exact arithmetic: the six upper entries are computed, the lower ones mirrored
#define M3_MIRROR(r) do { (r).m[3] = (r).m[1]; (r).m[6] = (r).m[2]; (r).m[7] = (r).m[5]; } while (0)
B2C_HD M3 m3_mul_sym(const M3& a, const M3& b) {
    M3 r;
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = i; j < 3; j++)
            r.m[i * 3 + j] = fma(a.m[i * 3 + 2], b.m[6 + j], fma(a.m[i * 3 + 1], b.m[3 + j], a.m[i * 3] * b.m[j]));
    M3_MIRROR(r);
    return r;
}
B2C_HD M3 m3_mul_tn_sym(const M3& a, const M3& b) {   // a' b
    M3 r;
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = i; j < 3; j++)
            r.m[i * 3 + j] = fma(a.m[6 + i], b.m[6 + j], fma(a.m[3 + i], b.m[3 + j], a.m[i] * b.m[j]));
    M3_MIRROR(r);
    return r;
}

#define DARE_MAX_ITER 33   // 2^33 steps: closed-loop |lambda| up to 1-1e-9; beyond that "no finite solution"

B2C_HD void dare_solve_one(int64_t i, int64_t B, int method, int guard, double dt,
       double q0, double q1, double q2, double r0, double r1,
       const double* __restrict__ v, const double* __restrict__ th,
       double* __restrict__ Kout, double* __restrict__ Pout,
       int32_t* __restrict__ iters, int32_t* __restrict__ status) {
    LinModel lm = b2c_linearize_point(b2c_guard_v(v[i], guard), th[i], dt, method);
    M3 A0 = {{lm.adiag, 0.0, lm.a13, 0.0, lm.adiag, lm.a23, 0.0, 0.0, lm.adiag}};
    // B (3x2): [[b11,b12],[b21,b22],[0,b32]]
    const double b11 = lm.b11, b12 = lm.b12, b21 = lm.b21, b22 = lm.b22, b32 = lm.b32;
    const double ir0 = 1.0 / r0, ir1 = 1.0 / r1;
    M3 A = A0, G, H;
    // G = B R^-1 B'
    G.m[0] = b11 * b11 * ir0 + b12 * b12 * ir1;
    G.m[1] = b11 * b21 * ir0 + b12 * b22 * ir1;
    G.m[2] = b12 * b32 * ir1;
    G.m[4] = b21 * b21 * ir0 + b22 * b22 * ir1;
    G.m[5] = b22 * b32 * ir1;
    G.m[8] = b32 * b32 * ir1;
    G.m[3] = G.m[1]; G.m[6] = G.m[2]; G.m[7] = G.m[5];
    H = {{q0, 0.0, 0.0, 0.0, q1, 0.0, 0.0, 0.0, q2}};
    int it = 0, st = B2C_MAXITER;
    for (; it < DARE_MAX_ITER; ++it) {
        M3 W = m3_mul(G, H);
        W.m[0] += 1.0; W.m[4] += 1.0; W.m[8] += 1.0;
        M3 Wi = m3_inv(W);
        M3 T = m3_mul(A, Wi);                 // A W^-1
        M3 HW = m3_mul_sym(H, Wi);                // H W^-1 = (I + H G)^-1 H: symmetric
        M3 Hn = m3_mul_tn_sym(A, m3_mul(HW, A));  // A' H W^-1 A: symmetric
        M3 Gn = m3_mul_sym(T, m3_mul_nt(G, A));   // A W^-1 G A': symmetric
        A = m3_mul(T, A);
        double dmax = 0.0, hmax = 0.0;
#pragma unroll
        for (int e = 0; e < 9; e++) {
            G.m[e] += Gn.m[e];
            H.m[e] += Hn.m[e];
        }
#pragma unroll
        for (int i = 0; i < 3; i++)
#pragma unroll
            for (int j = i; j < 3; j++) {
                dmax = fmax(dmax, fabs(Hn.m[i * 3 + j]));
                hmax = fmax(hmax, fabs(H.m[i * 3 + j]));
            }
        if (!(hmax < 1e300)) { st = B2C_NONFINITE; break; }   // catches NaN and overflow
        if (dmax <= 1e-15 * hmax) {
            // A_k = (closed loop)^(2^k) must have died out; if it has not, the pencil has an
            // eigenvalue on the unit circle (e.g. v = 0: scipy raises LinAlgError there)
            double amax = 0.0;
#pragma unroll
            for (int e = 0; e < 9; e++) amax = fmax(amax, fabs(A.m[e]));
            st = (amax < 1e-3) ? B2C_OK : B2C_MAXITER; ++it; break;
        }
    }
    // K = (R + B'PB)^-1 B'PA   (lqr_controller.py:130-132), P = H
    // PB (3x2)
    double pb[6], pa[9];
#pragma unroll
    for (int r = 0; r < 3; r++) {
        pb[r * 2 + 0] = H.m[r * 3] * b11 + H.m[r * 3 + 1] * b21;
        pb[r * 2 + 1] = H.m[r * 3] * b12 + H.m[r * 3 + 1] * b22 + H.m[r * 3 + 2] * b32;
    }
    M3 PA = m3_mul(H, A0);
#pragma unroll
    for (int e = 0; e < 9; e++) pa[e] = PA.m[e];
    double s00 = r0 + b11 * pb[0] + b21 * pb[2];
    double s01 = b11 * pb[1] + b21 * pb[3];
    double s10 = b12 * pb[0] + b22 * pb[2] + b32 * pb[4];
    double s11 = r1 + b12 * pb[1] + b22 * pb[3] + b32 * pb[5];
    double sym = 0.5 * (s01 + s10);
    double det = s00 * s11 - sym * sym;
    double idet = 1.0 / det;
    double kk[6];
#pragma unroll
    for (int c = 0; c < 3; c++) {
        double g0 = b11 * pa[c] + b21 * pa[3 + c];                       // (B'PA)[0][c]
        double g1 = b12 * pa[c] + b22 * pa[3 + c] + b32 * pa[6 + c];     // (B'PA)[1][c]
        kk[c] = (s11 * g0 - sym * g1) * idet;
        kk[3 + c] = (s00 * g1 - sym * g0) * idet;
    }
    bool fin = true;
#pragma unroll
    for (int e = 0; e < 6; e++) fin = fin && (fabs(kk[e]) < 1e300);
    if (!fin && st == B2C_OK) st = B2C_NONFINITE;
    if (st != B2C_OK) {           // fallback gain of compute_gain, lqr_controller.py:134-141
        kk[0] = 1.0; kk[1] = 0.0; kk[2] = 0.0; kk[3] = 0.0; kk[4] = 0.0; kk[5] = 1.0;
    }
#pragma unroll
    for (int e = 0; e < 6; e++) Kout[e * B + i] = kk[e];
    if (Pout) {
        Pout[0 * B + i] = H.m[0]; Pout[1 * B + i] = H.m[1]; Pout[2 * B + i] = H.m[2];
        Pout[3 * B + i] = H.m[4]; Pout[4 * B + i] = H.m[5]; Pout[5 * B + i] = H.m[8];
    }
    if (iters) iters[i] = it;
    if (status) status[i] = st;
}

__global__ void __launch_bounds__(128)
k_dare(int64_t B, int method, int guard, double dt,
       double q0, double q1, double q2, double r0, double r1,
       const double* __restrict__ v, const double* __restrict__ th,
       double* __restrict__ Kout, double* __restrict__ Pout,
       int32_t* __restrict__ iters, int32_t* __restrict__ status) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < B) dare_solve_one(i, B, method, guard, dt, q0, q1, q2, r0, r1, v, th, Kout, Pout, iters, status);
}

int b2c_dare_dev(b2c_ctx* ctx, int64_t B, int method, int guard, double dt,
                 const double Qd[3], const double Rd[2], const double* v, const double* th,
                 double* K, double* P, int32_t* iters, int32_t* status) {
    if (!ctx) { b2c_set_error("null context"); return -1; }
    if (method != B2C_LIN_EULER && method != B2C_LIN_EXACT) {
        b2c_set_error("DARE needs a discrete model (EULER or EXACT)"); return -2; }
    if (!(Rd[0] > 0.0) || !(Rd[1] > 0.0)) { b2c_set_error("R must be positive definite"); return -3; }
    if (B <= 0) return 0;
    B2C_DEVICE_GUARD(ctx);
    k_dare<<<b2c_blocks(B, 128), 128, 0, ctx->stream>>>(B, method, guard, dt, Qd[0], Qd[1], Qd[2],
                                                         Rd[0], Rd[1], v, th, K, P, iters, status);
    B2C_CHECK_LAUNCH(ctx);
    return 0;
}

// ---------------------------------------------------------------------------------------------
// LQR control law u = clip(u_ref - K wrap(x - x_ref))   (lqr_controller.py:149-189)
__global__ void __launch_bounds__(256)
k_lqr_control(int64_t B, double v_max, double w_max,
              const double* __restrict__ x, const double* __restrict__ xref,
              const double* __restrict__ uref, const double* __restrict__ K,
              double* __restrict__ u, double* __restrict__ err) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B) return;
    double e0 = x[i] - xref[i], e1 = x[B + i] - xref[B + i];
    double e2 = b2c_wrap(x[2 * B + i] - xref[2 * B + i]);
    // numpy evaluates -K @ e then adds u_ref: keep that association
    double f0 = -K[0 * B + i] * e0 + -K[1 * B + i] * e1 + -K[2 * B + i] * e2;
    double f1 = -K[3 * B + i] * e0 + -K[4 * B + i] * e1 + -K[5 * B + i] * e2;
    u[i] = b2c_clip(uref[i] + f0, -v_max, v_max);
    u[B + i] = b2c_clip(uref[B + i] + f1, -w_max, w_max);
    if (err) { err[i] = e0; err[B + i] = e1; err[2 * B + i] = e2; }
}

int b2c_lqr_control_dev(b2c_ctx* ctx, int64_t B, double v_max, double w_max,
                        const double* x, const double* xref, const double* uref,
                        const double* K, double* u, double* err) {
    if (!ctx) { b2c_set_error("null context"); return -1; }
    if (B <= 0) return 0;
    B2C_DEVICE_GUARD(ctx);
    k_lqr_control<<<b2c_blocks(B, 256), 256, 0, ctx->stream>>>(B, v_max, w_max, x, xref, uref, K, u, err);
    B2C_CHECK_LAUNCH(ctx);
    return 0;
}

// the same law on numpy-layout buffers: x, xref, err (B,3); uref, u (B,2); K (B,2,3)
__global__ void __launch_bounds__(256)
k_lqr_control_aos(int64_t B, double v_max, double w_max,
                  const double* __restrict__ x, const double* __restrict__ xref,
                  const double* __restrict__ uref, const double* __restrict__ K,
                  double* __restrict__ u, double* __restrict__ err) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B) return;
    const double* k = K + i * 6;
    double e0 = x[3 * i] - xref[3 * i], e1 = x[3 * i + 1] - xref[3 * i + 1];
    double e2 = b2c_wrap(x[3 * i + 2] - xref[3 * i + 2]);
    double f0 = -k[0] * e0 + -k[1] * e1 + -k[2] * e2;
    double f1 = -k[3] * e0 + -k[4] * e1 + -k[5] * e2;
    u[2 * i] = b2c_clip(uref[2 * i] + f0, -v_max, v_max);
    u[2 * i + 1] = b2c_clip(uref[2 * i + 1] + f1, -w_max, w_max);
    if (err) { err[3 * i] = e0; err[3 * i + 1] = e1; err[3 * i + 2] = e2; }
}

int b2c_lqr_control_host(b2c_ctx* ctx, int64_t B, double v_max, double w_max,
                         const double* x, const double* xref, const double* uref,
                         const double* K, double* u, double* err) {
    if (!ctx || !x || !xref || !uref || !K || !u) { b2c_set_error("null argument"); return -1; }
    if (B <= 0) return 0;
    B2C_DEVICE_GUARD(ctx);
    const size_t b3 = (size_t)B * 24, b2 = (size_t)B * 16, b6 = (size_t)B * 48;
    auto al = [](size_t v) { return (v + 255) / 256 * 256; };
    int rc = b2c_reserve(&ctx->dstage, &ctx->dstage_bytes, 3 * al(b3) + 2 * al(b2) + al(b6), false);
    if (rc) return rc;
    char* d = (char*)ctx->dstage;
    double* dx = (double*)d; d += al(b3);
    double* dxr = (double*)d; d += al(b3);
    double* de = (double*)d; d += al(b3);
    double* dur = (double*)d; d += al(b2);
    double* du = (double*)d; d += al(b2);
    double* dK = (double*)d;
    cudaStream_t s = ctx->stream;
    B2C_CUDA(cudaMemcpyAsync(dx, x, b3, cudaMemcpyHostToDevice, s));
    B2C_CUDA(cudaMemcpyAsync(dxr, xref, b3, cudaMemcpyHostToDevice, s));
    B2C_CUDA(cudaMemcpyAsync(dur, uref, b2, cudaMemcpyHostToDevice, s));
    B2C_CUDA(cudaMemcpyAsync(dK, K, b6, cudaMemcpyHostToDevice, s));
    k_lqr_control_aos<<<b2c_blocks(B, 256), 256, 0, s>>>(B, v_max, w_max, dx, dxr, dur, dK, du, err ? de : nullptr);
    B2C_CHECK_LAUNCH(ctx);
    B2C_CUDA(cudaMemcpyAsync(u, du, b2, cudaMemcpyDeviceToHost, s));
    if (err) B2C_CUDA(cudaMemcpyAsync(err, de, b3, cudaMemcpyDeviceToHost, s));
    B2C_CUDA(cudaStreamSynchronize(s));
    return 0;
}

#ifdef B2C_HOST_EMUL
// Test-only: the same per-instance code on the host (tests/ "not gpu" coverage of the kernel math).
extern "C" int b2c_emul_dare(int64_t B, int method, int guard, double dt, const double Qd[3],
                             const double Rd[2], const double* v, const double* th,
                             double* K, double* P, int32_t* iters, int32_t* status) {
    for (int64_t i = 0; i < B; ++i)
        dare_solve_one(i, B, method, guard, dt, Qd[0], Qd[1], Qd[2], Rd[0], Rd[1], v, th, K, P, iters, status);
    return 0;
}
extern "C" int b2c_emul_linearize(int64_t B, int method, int guard, double dt, const double* v,
                                  const double* th, double* Ad, double* Bd) {
    for (int64_t i = 0; i < B; ++i) {
        LinModel m = b2c_linearize_point(b2c_guard_v(v[i], guard), th[i], dt, method);
        double a[9] = {m.adiag, 0, m.a13, 0, m.adiag, m.a23, 0, 0, m.adiag};
        double b[6] = {m.b11, m.b12, m.b21, m.b22, 0, m.b32};
        for (int e = 0; e < 9; e++) Ad[e * B + i] = a[e];
        for (int e = 0; e < 6; e++) Bd[e * B + i] = b[e];
    }
    return 0;
}
#endif
