// K3: batched MPC QP.  One thread per instance; projected semismooth Newton whose linear solve is
// a 3-state / 2-input Riccati sweep held in registers; per-stage data in coalesced SoA scratch.
//
// Reference problem (HC/controllers/mpc_controller.py:345-522 solve_with_ltv, :150-314 solve):
//   min  sum_k [ (z_k-zc_k)'Q(z_k-zc_k) + (w_j+uo_k)'R(w_j+uo_k) ] + (z_N-zc_N)'P(z_N-zc_N) + rho*sum s^2
//   s.t. z_{k+1} = A_k z_k + B_k w_{j(k)},  j(k) = min(k / block_size, Nb-1)
//        -umax <= w_j + uo_k <= umax,   n_km.(p_k - o_m) >= d_safe + r_m - s_km,  s >= 0
//   LTV: z = x - x_ref (error), w = du, uo_k = u_ref,k, zc = 0      LTI: z = x, w = u, uo = 0, zc = x_ref
// Method: the slack of each soft row is eliminated in closed form,
//   min_{s>=0} rho s^2  s.t. s >= g   ==   rho * max(0, g)^2 ,
// which leaves a strictly convex, C^1, piecewise-quadratic optimal-control problem with bounds on
// the inputs only.  Each Newton iteration = backward Riccati sweep on the current active sets
// (violated rows add rho*n n' to the stage Hessian; inputs held at a bound by the sign of the true
// gradient are frozen) + a forward roll-out with feedback, clipping and a monotone line search.
// When the active sets stop changing the step lands on the exact minimiser (to rounding), i.e. the
// same point CVXPY/OSQP approximates to 1e-5; typical count is 1-8 iterations.
#include <string.h>
#include "mpc_args.cuh"

namespace {

// scratch field map (each field is a length-B column; instance i lives at field*B + i)
template <bool BS1> struct Lay {
    static constexpr int MODEL = BS1 ? 8 : 28;
    __host__ __device__ static int ab(int k) { return 4 * k; }                       // a,b,c,s
    __host__ __device__ static int w(int N, int Nb, int buf, int j) { return 4 * N + buf * 2 * Nb + 2 * j; }
    __host__ __device__ static int model(int N, int Nb, int j) { return 4 * N + 4 * Nb + MODEL * j; }
    __host__ __device__ static int gain(int N, int Nb, int j) { return 4 * N + 4 * Nb + MODEL * Nb + 8 * j; }
    __host__ __device__ static int total(int N, int Nb) { return 4 * N + 4 * Nb + (MODEL + 8) * Nb; }
};

#define SC(f) a.scratch[(int64_t)(f) * a.B + i]

// ----- forward pass ---------------------------------------------------------------------------
// Rolls the plant out from z0 under either the current inputs (USE_GAIN=false) or the Newton
// update  w+ = clip(w - alpha*kf - K (z+ - z))  (USE_GAIN=true, written to buffer `dst`), and
// accumulates cost + the per-super-stage quadratic model at the NEW trajectory.
//
// Hard rows (use_soft_constraints=False, :197-200 / :260-270 / :383-386 / :464-468) are handled by the method of
// multipliers on the same machinery: with multiplier lm >= 0 and penalty mu a row contributes
//   (mu/2) * max(0, g + lm/mu)^2      (g = d_safe + r - n.(p - o), the row's violation),
// i.e. exactly the soft row with rho -> mu/2 and a shifted threshold.  `mu` = 0 selects the soft formula;
// `update` applies lm <- max(0, lm + mu*g) in place before evaluating.  In hard mode `max_viol` returns the KKT
// residual max |max(g, -lm/mu)| (primal violation and complementarity) and `J0` the objective without row terms.
template <bool BS1, bool USE_GAIN>
B2C_HD double mpc_forward(const MpcArgs& a, int64_t i, double z0x, double z0y, double z0t,
                                              int src, int dst, double alpha, double& max_viol,
                                              bool& infeasible, double& lamx, double& lamy, double& lamt,
                                              const double mu = 0.0, const bool update = false, double* J0out = nullptr) {
    using L = Lay<BS1>;
    const int N = a.N, Nb = a.Nb, M = a.M;
    const bool ltv = (a.kind == B2C_MPC_LTV);
    double zx = z0x, zy = z0y, zt = z0t;          // new trajectory
    double ox_ = z0x, oy_ = z0y, ot_ = z0t;       // old trajectory (only with gains)
    double J = 0.0, Jrows = 0.0;
    const bool hard = mu > 0.0;
    const double rho = hard ? 0.5 * mu : a.rho;
    max_viol = 0.0;
    for (int j = 0; j < Nb; ++j) {
        const int k_begin = j * a.bs;
        const int k_end = (j == Nb - 1) ? N : (k_begin + a.bs);
        // bounds of this block: intersection over its steps of [-umax - uo_k, umax - uo_k]
        double lb0 = -a.v_max, ub0 = a.v_max, lb1 = -a.w_max, ub1 = a.w_max;
        if (ltv) {
            lb0 = -1e300; ub0 = 1e300; lb1 = -1e300; ub1 = 1e300;
            for (int k = k_begin; k < k_end; ++k) {
                double ur0 = B2C_LDG(a.uref + (int64_t)(2 * k) * a.B + i);
                double ur1 = B2C_LDG(a.uref + (int64_t)(2 * k + 1) * a.B + i);
                lb0 = fmax(lb0, -a.v_max - ur0); ub0 = fmin(ub0, a.v_max - ur0);
                lb1 = fmax(lb1, -a.w_max - ur1); ub1 = fmin(ub1, a.w_max - ur1);
            }
        }
        if (lb0 > ub0 || lb1 > ub1) infeasible = true;
        double w0, w1, wo0 = 0.0, wo1 = 0.0;
        if (USE_GAIN) {
            wo0 = SC(L::w(N, Nb, src, j)); wo1 = SC(L::w(N, Nb, src, j) + 1);
            const int g = L::gain(N, Nb, j);
            double dx = zx - ox_, dy = zy - oy_, dt_ = zt - ot_;
            w0 = wo0 - alpha * SC(g + 6) - (SC(g + 0) * dx + SC(g + 1) * dy + SC(g + 2) * dt_);
            w1 = wo1 - alpha * SC(g + 7) - (SC(g + 3) * dx + SC(g + 4) * dy + SC(g + 5) * dt_);
            w0 = b2c_clip(w0, lb0, ub0); w1 = b2c_clip(w1, lb1, ub1);
            SC(L::w(N, Nb, dst, j)) = w0; SC(L::w(N, Nb, dst, j) + 1) = w1;
        } else {
            w0 = SC(L::w(N, Nb, src, j)); w1 = SC(L::w(N, Nb, src, j) + 1);
        }
        // block accumulators (only the general path needs them)
        double al = 0.0, be = 0.0;                             // Phi = I + [0 0 al; 0 0 be; 0 0 0]
        double g00 = 0.0, g01 = 0.0, g10 = 0.0, g11 = 0.0, g21 = 0.0;   // Gamma (3x2), Gamma[2][0] == 0
        double Hzz[6] = {0, 0, 0, 0, 0, 0};                    // 00 01 02 11 12 22
        double Hzu[6] = {0, 0, 0, 0, 0, 0};                    // row-major 3x2
        double Huu[3] = {0, 0, 0};                             // 00 01 11
        double gz[3] = {0, 0, 0}, gu[2] = {0, 0};
        for (int k = k_begin; k < k_end; ++k) {
            double ex = zx, ey = zy, et = zt, uo0 = 0.0, uo1 = 0.0;
            const double prx = B2C_LDG(a.xref + (int64_t)(3 * k) * a.B + i);
            const double pry = B2C_LDG(a.xref + (int64_t)(3 * k + 1) * a.B + i);
            if (ltv) {
                uo0 = B2C_LDG(a.uref + (int64_t)(2 * k) * a.B + i);
                uo1 = B2C_LDG(a.uref + (int64_t)(2 * k + 1) * a.B + i);
            } else {
                ex -= prx; ey -= pry; et -= B2C_LDG(a.xref + (int64_t)(3 * k + 2) * a.B + i);
            }
            double wxx = a.q0, wxy = 0.0, wyy = a.q1;
            double gx = a.q0 * ex, gy = a.q1 * ey, gt = a.q2 * et;
            J += a.q0 * ex * ex + a.q1 * ey * ey + a.q2 * et * et;
            for (int m = 0; m < M; ++m) {                      // soft obstacle rows, :439-468 / :238-270
                const double obx = B2C_LDG(a.obs + (int64_t)(3 * m) * a.B + i);
                const double oby = B2C_LDG(a.obs + (int64_t)(3 * m + 1) * a.B + i);
                const double obr = B2C_LDG(a.obs + (int64_t)(3 * m + 2) * a.B + i);
                const double ddx = prx - obx, ddy = pry - oby;
                const double d2 = ddx * ddx + ddy * ddy;
                if (d2 > 1e-4) {                               // dist > 0.01
                    const double inv = B2C_RSQRT(d2);
                    const double nx = ddx * inv, ny = ddy * inv;
                    // LTV: p = p_ref + dz  => n.(p - o) = dist + n.dz ;  LTI: p = z
                    double lhs = ltv ? (d2 * inv + nx * zx + ny * zy) : (nx * (zx - obx) + ny * (zy - oby));
                    double viol = (a.d_safe + obr) - lhs;
                    if (hard) {
                        if (k == 0 && viol > 1e-6) infeasible = true;      // a row on the fixed initial state
                        const int lf = a.lam_off + k * M + m;
                        double lm = SC(lf);
                        max_viol = fmax(max_viol, fabs(fmax(viol, -lm / mu)));
                        if (update) { lm = fmax(0.0, lm + mu * viol); SC(lf) = lm; }
                        viol += lm / mu;
                    } else if (viol > 0.0) {
                        max_viol = fmax(max_viol, viol);
                    }
                    if (viol > 0.0) {
                        Jrows += rho * viol * viol;
                        const double rn0 = rho * nx, rn1 = rho * ny;
                        wxx += rn0 * nx; wxy += rn0 * ny; wyy += rn1 * ny;
                        gx -= rn0 * viol; gy -= rn1 * viol;
                    }
                }
            }
            const double ut0 = w0 + uo0, ut1 = w1 + uo1;
            J += a.r0 * ut0 * ut0 + a.r1 * ut1 * ut1;
            const double a_k = SC(L::ab(k)), b_k = SC(L::ab(k) + 1), c_k = SC(L::ab(k) + 2), s_k = SC(L::ab(k) + 3);
            if (BS1) {
                const int mo = L::model(N, Nb, j);
                SC(mo + 0) = gx; SC(mo + 1) = gy; SC(mo + 2) = gt;
                SC(mo + 3) = wxx; SC(mo + 4) = wxy; SC(mo + 5) = wyy;
                SC(mo + 6) = a.r0 * ut0; SC(mo + 7) = a.r1 * ut1;
            } else {
                const double t0 = wxx * al + wxy * be, t1 = wxy * al + wyy * be;
                Hzz[0] += wxx; Hzz[1] += wxy; Hzz[2] += t0; Hzz[3] += wyy; Hzz[4] += t1;
                Hzz[5] += al * t0 + be * t1 + a.q2;
                const double wg00 = wxx * g00 + wxy * g10, wg01 = wxx * g01 + wxy * g11;
                const double wg10 = wxy * g00 + wyy * g10, wg11 = wxy * g01 + wyy * g11;
                const double wg21 = a.q2 * g21;
                Hzu[0] += wg00; Hzu[1] += wg01; Hzu[2] += wg10; Hzu[3] += wg11;
                Hzu[4] += al * wg00 + be * wg10; Hzu[5] += al * wg01 + be * wg11 + wg21;
                Huu[0] += g00 * wg00 + g10 * wg10 + a.r0;
                Huu[1] += g00 * wg01 + g10 * wg11;
                Huu[2] += g01 * wg01 + g11 * wg11 + g21 * wg21 + a.r1;
                gz[0] += gx; gz[1] += gy; gz[2] += al * gx + be * gy + gt;
                gu[0] += g00 * gx + g10 * gy + a.r0 * ut0;
                gu[1] += g01 * gx + g11 * gy + g21 * gt + a.r1 * ut1;
                // Gamma <- A_k Gamma + B_k ; Phi <- A_k Phi
                g01 += a_k * g21; g11 += b_k * g21;
                g00 += c_k; g10 += s_k; g21 += a.dt;
                al += a_k; be += b_k;
            }
            // advance plant(s)
            const double nzx = zx + a_k * zt + c_k * w0, nzy = zy + b_k * zt + s_k * w0;
            zt = zt + a.dt * w1; zx = nzx; zy = nzy;
            if (USE_GAIN) {
                const double nox = ox_ + a_k * ot_ + c_k * wo0, noy = oy_ + b_k * ot_ + s_k * wo0;
                ot_ = ot_ + a.dt * wo1; ox_ = nox; oy_ = noy;
            }
        }
        if (!BS1) {
            const int mo = L::model(N, Nb, j);
#pragma unroll
            for (int e = 0; e < 6; e++) SC(mo + e) = Hzz[e];
#pragma unroll
            for (int e = 0; e < 6; e++) SC(mo + 6 + e) = Hzu[e];
            SC(mo + 12) = Huu[0]; SC(mo + 13) = Huu[1]; SC(mo + 14) = Huu[2];
            SC(mo + 15) = gz[0]; SC(mo + 16) = gz[1]; SC(mo + 17) = gz[2];
            SC(mo + 18) = gu[0]; SC(mo + 19) = gu[1];
            SC(mo + 20) = al; SC(mo + 21) = be;
            SC(mo + 22) = g00; SC(mo + 23) = g01; SC(mo + 24) = g10; SC(mo + 25) = g11; SC(mo + 26) = g21;
        }
    }
    // terminal cost (mpc_controller.py:412 / :212-213)
    double ex = zx, ey = zy, et = zt;
    if (!ltv) {
        ex -= B2C_LDG(a.xref + (int64_t)(3 * N) * a.B + i);
        ey -= B2C_LDG(a.xref + (int64_t)(3 * N + 1) * a.B + i);
        et -= B2C_LDG(a.xref + (int64_t)(3 * N + 2) * a.B + i);
    }
    J += a.p0 * ex * ex + a.p1 * ey * ey + a.p2 * et * et;
    lamx = a.p0 * ex; lamy = a.p1 * ey; lamt = a.p2 * et;
    if (J0out) *J0out = J;
    return J + Jrows;
}

// ----- backward pass --------------------------------------------------------------------------
// Riccati sweep on the stored model.  Returns the Newton decrement 1/2 sum qu' Quu^-1 qu.
template <bool BS1>
B2C_HD double mpc_backward(const MpcArgs& a, int64_t i, int cur,
                                               double lamx, double lamy, double lamt) {
    using L = Lay<BS1>;
    const int N = a.N, Nb = a.Nb;
    const bool ltv = (a.kind == B2C_MPC_LTV);
    // value function 1/2 dz'S dz + sv'dz  and the true costate lam
    double s00 = a.p0, s01 = 0.0, s02 = 0.0, s11 = a.p1, s12 = 0.0, s22 = a.p2;
    double v0 = lamx, v1 = lamy, v2 = lamt;
    double l0 = lamx, l1 = lamy, l2 = lamt;
    double dec = 0.0;
    for (int j = Nb - 1; j >= 0; --j) {
        const int mo = L::model(N, Nb, j);
        double Hzz0, Hzz1, Hzz2, Hzz3, Hzz4, Hzz5, Hzu0, Hzu1, Hzu2, Hzu3, Hzu4, Hzu5, Huu0, Huu1, Huu2;
        double gz0, gz1, gz2, gu0, gu1, al, be, g00, g01, g10, g11, g21;
        if (BS1) {
            gz0 = SC(mo + 0); gz1 = SC(mo + 1); gz2 = SC(mo + 2);
            Hzz0 = SC(mo + 3); Hzz1 = SC(mo + 4); Hzz3 = SC(mo + 5);
            Hzz2 = 0.0; Hzz4 = 0.0; Hzz5 = a.q2;
            Hzu0 = Hzu1 = Hzu2 = Hzu3 = Hzu4 = Hzu5 = 0.0;
            Huu0 = a.r0; Huu1 = 0.0; Huu2 = a.r1;
            gu0 = SC(mo + 6); gu1 = SC(mo + 7);
            al = SC(L::ab(j)); be = SC(L::ab(j) + 1);
            g00 = SC(L::ab(j) + 2); g10 = SC(L::ab(j) + 3); g01 = 0.0; g11 = 0.0; g21 = a.dt;
        } else {
            Hzz0 = SC(mo + 0); Hzz1 = SC(mo + 1); Hzz2 = SC(mo + 2); Hzz3 = SC(mo + 3); Hzz4 = SC(mo + 4); Hzz5 = SC(mo + 5);
            Hzu0 = SC(mo + 6); Hzu1 = SC(mo + 7); Hzu2 = SC(mo + 8); Hzu3 = SC(mo + 9); Hzu4 = SC(mo + 10); Hzu5 = SC(mo + 11);
            Huu0 = SC(mo + 12); Huu1 = SC(mo + 13); Huu2 = SC(mo + 14);
            gz0 = SC(mo + 15); gz1 = SC(mo + 16); gz2 = SC(mo + 17);
            gu0 = SC(mo + 18); gu1 = SC(mo + 19);
            al = SC(mo + 20); be = SC(mo + 21);
            g00 = SC(mo + 22); g01 = SC(mo + 23); g10 = SC(mo + 24); g11 = SC(mo + 25); g21 = SC(mo + 26);
        }
        // true gradient wrt w_j (decides which bounded inputs are frozen), then costate update
        const double gt0 = gu0 + g00 * l0 + g10 * l1;
        const double gt1 = gu1 + g01 * l0 + g11 * l1 + g21 * l2;
        {
            const double nl2 = gz2 + al * l0 + be * l1 + l2;
            l0 = gz0 + l0; l1 = gz1 + l1; l2 = nl2;
        }
        // bounds of the block (same rule as the forward pass)
        double lb0 = -a.v_max, ub0 = a.v_max, lb1 = -a.w_max, ub1 = a.w_max;
        if (ltv) {
            const int k_begin = j * a.bs, k_end = (j == Nb - 1) ? N : (k_begin + a.bs);
            lb0 = -1e300; ub0 = 1e300; lb1 = -1e300; ub1 = 1e300;
            for (int k = k_begin; k < k_end; ++k) {
                double ur0 = B2C_LDG(a.uref + (int64_t)(2 * k) * a.B + i);
                double ur1 = B2C_LDG(a.uref + (int64_t)(2 * k + 1) * a.B + i);
                lb0 = fmax(lb0, -a.v_max - ur0); ub0 = fmin(ub0, a.v_max - ur0);
                lb1 = fmax(lb1, -a.w_max - ur1); ub1 = fmin(ub1, a.w_max - ur1);
            }
        }
        const double w0 = SC(L::w(N, Nb, cur, j)), w1 = SC(L::w(N, Nb, cur, j) + 1);
        const bool c0 = (w0 <= lb0 && gt0 > 0.0) || (w0 >= ub0 && gt0 < 0.0);
        const bool c1 = (w1 <= lb1 && gt1 > 0.0) || (w1 >= ub1 && gt1 < 0.0);
        // SA = S * Abar : first two columns are S's, third is t
        const double t0 = al * s00 + be * s01 + s02;
        const double t1 = al * s01 + be * s11 + s12;
        const double t2 = al * s02 + be * s12 + s22;
        // Qzz = Hzz + Abar' S Abar
        const double Q00 = Hzz0 + s00, Q01 = Hzz1 + s01, Q02 = Hzz2 + t0;
        const double Q11 = Hzz3 + s11, Q12 = Hzz4 + t1;
        const double Q22 = Hzz5 + al * t0 + be * t1 + t2;
        // Quz = Hzu' + Bbar' S Abar   (2x3)
        double U00 = Hzu0 + g00 * s00 + g10 * s01;
        double U01 = Hzu2 + g00 * s01 + g10 * s11;
        double U02 = Hzu4 + g00 * t0 + g10 * t1;
        double U10 = Hzu1 + g01 * s00 + g11 * s01 + g21 * s02;
        double U11 = Hzu3 + g01 * s01 + g11 * s11 + g21 * s12;
        double U12 = Hzu5 + g01 * t0 + g11 * t1 + g21 * t2;
        // Quu = Huu + Bbar' S Bbar
        const double sb00 = s00 * g00 + s01 * g10, sb10 = s01 * g00 + s11 * g10, sb20 = s02 * g00 + s12 * g10;
        const double sb01 = s00 * g01 + s01 * g11 + s02 * g21;
        const double sb11 = s01 * g01 + s11 * g11 + s12 * g21;
        const double sb21 = s02 * g01 + s12 * g11 + s22 * g21;
        double R00 = Huu0 + g00 * sb00 + g10 * sb10;
        double R01 = Huu1 + g00 * sb01 + g10 * sb11;
        double R11 = Huu2 + g01 * sb01 + g11 * sb11 + g21 * sb21;
        (void)sb20;
        // qz, qu
        const double qz0 = gz0 + v0, qz1 = gz1 + v1, qz2 = gz2 + al * v0 + be * v1 + v2;
        double qu0 = gu0 + g00 * v0 + g10 * v1;
        double qu1 = gu1 + g01 * v0 + g11 * v1 + g21 * v2;
        if (c0) { R00 = 1.0; R01 = 0.0; U00 = 0.0; U01 = 0.0; U02 = 0.0; qu0 = 0.0; }
        if (c1) { R11 = 1.0; R01 = 0.0; U10 = 0.0; U11 = 0.0; U12 = 0.0; qu1 = 0.0; }
        const double idet = 1.0 / (R00 * R11 - R01 * R01);
        const double i00 = R11 * idet, i01 = -R01 * idet, i11 = R00 * idet;
        const double K00 = i00 * U00 + i01 * U10, K01 = i00 * U01 + i01 * U11, K02 = i00 * U02 + i01 * U12;
        const double K10 = i01 * U00 + i11 * U10, K11 = i01 * U01 + i11 * U11, K12 = i01 * U02 + i11 * U12;
        const double k0 = i00 * qu0 + i01 * qu1, k1 = i01 * qu0 + i11 * qu1;
        const int g = L::gain(N, Nb, j);
        SC(g + 0) = K00; SC(g + 1) = K01; SC(g + 2) = K02; SC(g + 3) = K10; SC(g + 4) = K11; SC(g + 5) = K12;
        SC(g + 6) = k0; SC(g + 7) = k1;
        dec += 0.5 * (qu0 * k0 + qu1 * k1);
        // S = Qzz - Quz' K ; sv = qz - Quz' kf
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
    return dec;
}

// ----- globalisation fallback -------------------------------------------------------------------
// Clipped closed-loop roll-outs are the fast path but carry no descent guarantee when many inputs
// saturate.  If they fail, the gains are rewritten in place as the OPEN-LOOP Newton direction of the
// linear model (d_j = -kf_j - K_j dz_j, dz_{j+1} = Abar dz_j + Bbar d_j), after which the forward pass
// evaluates the projection arc P[w + alpha d] -- Bertsekas' projected Newton method, for which a
// small enough alpha always decreases the cost because the frozen inputs are decoupled in Quu.
template <bool BS1>
B2C_HD void mpc_gains_to_open_loop(const MpcArgs& a, int64_t i) {
    using L = Lay<BS1>;
    const int N = a.N, Nb = a.Nb;
    double dx = 0.0, dy = 0.0, dth = 0.0;
    for (int j = 0; j < Nb; ++j) {
        const int g = L::gain(N, Nb, j);
        const double d0 = -SC(g + 6) - (SC(g + 0) * dx + SC(g + 1) * dy + SC(g + 2) * dth);
        const double d1 = -SC(g + 7) - (SC(g + 3) * dx + SC(g + 4) * dy + SC(g + 5) * dth);
        double al, be, g00, g01, g10, g11, g21;
        if (BS1) {
            al = SC(L::ab(j)); be = SC(L::ab(j) + 1); g00 = SC(L::ab(j) + 2); g10 = SC(L::ab(j) + 3);
            g01 = 0.0; g11 = 0.0; g21 = a.dt;
        } else {
            const int mo = L::model(N, Nb, j);
            al = SC(mo + 20); be = SC(mo + 21);
            g00 = SC(mo + 22); g01 = SC(mo + 23); g10 = SC(mo + 24); g11 = SC(mo + 25); g21 = SC(mo + 26);
        }
        const double ndx = dx + al * dth + g00 * d0 + g01 * d1;
        const double ndy = dy + be * dth + g10 * d0 + g11 * d1;
        dth = dth + g21 * d1; dx = ndx; dy = ndy;
#pragma unroll
        for (int e = 0; e < 6; e++) SC(g + e) = 0.0;
        SC(g + 6) = -d0; SC(g + 7) = -d1;
    }
}

// ----- the solver kernel -----------------------------------------------------------------------
#define MPC_TPB 128

template <bool BS1>
B2C_HD void mpc_solve_one(const MpcArgs& a, const int64_t i) {
    using L = Lay<BS1>;
    const int N = a.N, Nb = a.Nb;
    const bool ltv = (a.kind == B2C_MPC_LTV);

    // ---- setup: linearize along the reference (K1 fused), initial state, initial inputs -------
    const double x0x = B2C_LDG(a.x0 + i), x0y = B2C_LDG(a.x0 + a.B + i), x0t = B2C_LDG(a.x0 + 2 * a.B + i);
    const double r0x = B2C_LDG(a.xref + i), r0y = B2C_LDG(a.xref + a.B + i), r0t = B2C_LDG(a.xref + 2 * a.B + i);
    double z0x, z0y, z0t;
    if (ltv) {                                    // mpc_controller.py:392-401, :421
        z0x = x0x - r0x; z0y = x0y - r0y; z0t = b2c_wrap(x0t - r0t);
        double th_prev = r0t, cum = 0.0;          // np.unwrap along the horizon
        for (int k = 0; k < N; ++k) {
            const double th_raw = B2C_LDG(a.xref + (int64_t)(3 * k + 2) * a.B + i);
            if (k > 0) cum += unwrap_correction(th_raw - th_prev);
            th_prev = th_raw;
            const double th = th_raw + cum;
            const double vr = b2c_guard_v(B2C_LDG(a.uref + (int64_t)(2 * k) * a.B + i), B2C_GUARD_MPC);   // :425
            double s, c; sincos(th, &s, &c);
            SC(L::ab(k) + 0) = -vr * s * a.dt; SC(L::ab(k) + 1) = vr * c * a.dt;
            SC(L::ab(k) + 2) = c * a.dt;       SC(L::ab(k) + 3) = s * a.dt;
        }
        for (int j = 0; j < Nb; ++j) { SC(L::w(N, Nb, 0, j)) = 0.0; SC(L::w(N, Nb, 0, j) + 1) = 0.0; }
    } else {                                      // :186-190, :223
        z0x = x0x; z0y = x0y; z0t = x0t;
        const double vr = b2c_guard_v(B2C_LDG(a.uref + i), B2C_GUARD_MPC);
        double s, c; sincos(r0t, &s, &c);
        for (int k = 0; k < N; ++k) {
            SC(L::ab(k) + 0) = -vr * s * a.dt; SC(L::ab(k) + 1) = vr * c * a.dt;
            SC(L::ab(k) + 2) = c * a.dt;       SC(L::ab(k) + 3) = s * a.dt;
            SC(L::w(N, Nb, 0, k)) = B2C_LDG(a.uref + (int64_t)(2 * k) * a.B + i);
            SC(L::w(N, Nb, 0, k) + 1) = B2C_LDG(a.uref + (int64_t)(2 * k + 1) * a.B + i);
        }
    }
    // clip the initial inputs into their boxes (bounds recomputed exactly as in the passes)
    bool infeasible = false;
    for (int j = 0; j < Nb; ++j) {
        double lb0 = -a.v_max, ub0 = a.v_max, lb1 = -a.w_max, ub1 = a.w_max;
        if (ltv) {
            const int k_begin = j * a.bs, k_end = (j == Nb - 1) ? N : (k_begin + a.bs);
            lb0 = -1e300; ub0 = 1e300; lb1 = -1e300; ub1 = 1e300;
            for (int k = k_begin; k < k_end; ++k) {
                double ur0 = B2C_LDG(a.uref + (int64_t)(2 * k) * a.B + i);
                double ur1 = B2C_LDG(a.uref + (int64_t)(2 * k + 1) * a.B + i);
                lb0 = fmax(lb0, -a.v_max - ur0); ub0 = fmin(ub0, a.v_max - ur0);
                lb1 = fmax(lb1, -a.w_max - ur1); ub1 = fmin(ub1, a.w_max - ur1);
            }
        }
        if (lb0 > ub0 || lb1 > ub1) infeasible = true;
        SC(L::w(N, Nb, 0, j)) = b2c_clip(SC(L::w(N, Nb, 0, j)), lb0, ub0);
        SC(L::w(N, Nb, 0, j) + 1) = b2c_clip(SC(L::w(N, Nb, 0, j) + 1), lb1, ub1);
    }

    // ---- Newton loop (state machine so that a warp's lanes share each pass) --------------------
    // hard rows: method of multipliers around it -- lm = 0, mu = 2 rho (the soft problem) to start with
    const bool hard = !a.soft && a.M > 0;
    double mu = hard ? 2.0 * a.rho : 0.0, J0 = 0.0, res_prev = 1e300;
    int outer = 0, stagnant = 0, softened = 0;
    if (hard) for (int e = 0; e < N * a.M; ++e) SC(a.lam_off + e) = 0.0;
    int cur = 0, it = 0, st = B2C_MAXITER;
    double lamx, lamy, lamt, max_viol = 0.0;
    double J = mpc_forward<BS1, false>(a, i, z0x, z0y, z0t, cur, cur, 0.0, max_viol, infeasible, lamx, lamy, lamt,
                                       mu, false, &J0);
    bool active = true;
    if (infeasible) { st = B2C_INFEASIBLE; active = false; }
    if (!(fabs(J) < 1e300)) { st = B2C_NONFINITE; active = false; }
    bool need_backward = true, force_outer = false;
    double alpha = 1.0, dec = 0.0;
    int ls = 0;
    while (active) {
        if (need_backward) {
            dec = mpc_backward<BS1>(a, i, cur, lamx, lamy, lamt);
            if (!(dec < 1e300)) { st = B2C_NONFINITE; break; }
            if (dec <= a.tol * fmax(1.0, fabs(J)) || force_outer) {
                force_outer = false;
                if (!hard || max_viol <= MPC_HARD_TOL) { st = B2C_OK; break; }
                // inner problem solved, rows not yet feasible / complementary: multiplier step.  A residual
                // that does not drop fourfold asks for a stiffer penalty; no progress at the stiffest means the
                // rows cannot be met (the reference's solver reports infeasible -> fallback)
                stagnant = (max_viol > 0.9 * res_prev) ? stagnant + 1 : 0;
                if (stagnant >= 2 && max_viol <= 1e-7) { st = B2C_OK; break; }      // at the rounding floor of the inner solves
                if (++outer > MPC_HARD_OUTER || (mu >= MPC_HARD_MU_MAX && stagnant >= 5)) { st = B2C_INFEASIBLE; break; }
                const bool stiffer = max_viol > 0.25 * res_prev && mu < MPC_HARD_MU_MAX;
                res_prev = max_viol;
                bool dummy = false;
                J = mpc_forward<BS1, false>(a, i, z0x, z0y, z0t, cur, cur, 0.0, max_viol, dummy, lamx, lamy, lamt,
                                            mu, true, &J0);
                if (stiffer) {
                    mu *= 10.0;
                    J = mpc_forward<BS1, false>(a, i, z0x, z0y, z0t, cur, cur, 0.0, max_viol, dummy, lamx, lamy, lamt,
                                                mu, false, &J0);
                }
                if (!(fabs(J) < 1e300)) { st = B2C_NONFINITE; break; }
                continue;
            }
            if (it >= a.max_iter) { st = B2C_MAXITER; break; }
            alpha = (it == 0) ? 1.0 : fmin(1.0, MPC_LS_GROW * alpha);
            ++it; ls = 0; need_backward = false;
        }
        double tlx, tly, tlt, tviol;
        bool dummy = false;
        double tJ0 = 0.0;
        const double Jn = mpc_forward<BS1, true>(a, i, z0x, z0y, z0t, cur, cur ^ 1, alpha, tviol, dummy, tlx, tly, tlt,
                                                 mu, false, &tJ0);
        if (Jn < J) {
            J = Jn; J0 = tJ0; cur ^= 1; lamx = tlx; lamy = tly; lamt = tlt; max_viol = tviol; need_backward = true;
        } else {
            alpha *= 0.5;
            ++ls;
            if (ls == MPC_LS_CLOSED) mpc_gains_to_open_loop<BS1>(a, i);     // the arc continues from this alpha
            if (ls >= MPC_LS_MAX) {               // cannot decrease any further: rounding floor or stall
                if (hard && dec > 1e-9 * fmax(1.0, fabs(J)) && mu > 2.0 * a.rho && softened < 8) {
                    // the penalty has become too stiff for the active-set Newton step: back off (any mu is a
                    // valid method-of-multipliers step; the multipliers are kept) and carry on from this point
                    mu *= 0.1; ++softened; res_prev = 1e300;
                    bool dm = false;
                    J = mpc_forward<BS1, false>(a, i, z0x, z0y, z0t, cur, cur, 0.0, max_viol, dm, lamx, lamy, lamt,
                                                mu, false, &J0);
                    need_backward = true;
                    continue;
                }
                st = (dec <= 1e-9 * fmax(1.0, fabs(J))) ? B2C_OK : B2C_MAXITER;
                if (st == B2C_OK && hard && max_viol > MPC_HARD_TOL) { need_backward = true; dec = 0.0; st = B2C_MAXITER; force_outer = true; continue; }
                break;
            }
        }
    }

    // ---- outputs ---------------------------------------------------------------------------------
    const double ur00 = B2C_LDG(a.uref + i), ur01 = B2C_LDG(a.uref + a.B + i);
    double u00, u01, Jout;
    int sc = a.step_count ? a.step_count[i] : a.ramp_up_steps;
    if (st == B2C_OK) {
        const double w0 = SC(L::w(N, Nb, cur, 0)), w1 = SC(L::w(N, Nb, cur, 0) + 1);
        u00 = ltv ? (ur00 + w0) : w0;
        u01 = ltv ? (ur01 + w1) : w1;
        Jout = hard ? J0 : J;                    // problem.value: no slack term in the hard problem
        if (hard) max_viol = 0.0;                // slack_used is False without slack variables, :294-296
        if (a.useq || a.xpred) {                 // :484-498: un-block, add the (wrapped) references
            double zx = z0x, zy = z0y, zt = z0t;
            for (int k = 0; k <= N; ++k) {
                if (a.xpred) {
                    double ax = zx, ay = zy, at = zt;
                    if (ltv) {
                        ax += B2C_LDG(a.xref + (int64_t)(3 * k) * a.B + i);
                        ay += B2C_LDG(a.xref + (int64_t)(3 * k + 1) * a.B + i);
                        at += B2C_LDG(a.xref + (int64_t)(3 * k + 2) * a.B + i);
                    }
                    a.xpred[(int64_t)(3 * k) * a.B + i] = ax;
                    a.xpred[(int64_t)(3 * k + 1) * a.B + i] = ay;
                    a.xpred[(int64_t)(3 * k + 2) * a.B + i] = at;
                }
                if (k == N) break;
                int j = k / a.bs; if (j > Nb - 1) j = Nb - 1;
                const double wk0 = SC(L::w(N, Nb, cur, j)), wk1 = SC(L::w(N, Nb, cur, j) + 1);
                if (a.useq) {
                    a.useq[(int64_t)(2 * k) * a.B + i] = ltv ? (wk0 + B2C_LDG(a.uref + (int64_t)(2 * k) * a.B + i)) : wk0;
                    a.useq[(int64_t)(2 * k + 1) * a.B + i] = ltv ? (wk1 + B2C_LDG(a.uref + (int64_t)(2 * k + 1) * a.B + i)) : wk1;
                }
                const double a_k = SC(L::ab(k)), b_k = SC(L::ab(k) + 1), c_k = SC(L::ab(k) + 2), s_k = SC(L::ab(k) + 3);
                const double nzx = zx + a_k * zt + c_k * wk0, nzy = zy + b_k * zt + s_k * wk0;
                zt = zt + a.dt * wk1; zx = nzx; zy = nzy;
            }
        }
        if (ltv) {                                // cold-start ramp, :500-507
            if (sc < a.ramp_up_steps) {
                const double lim = a.w_max * ((double)(sc + 1) / (double)a.ramp_up_steps);
                u01 = b2c_clip(u01, -lim, lim);
                if (a.useq) a.useq[a.B + i] = u01;
            }
            if (a.step_count) a.step_count[i] = sc + 1;
        }
    } else {                                      // _get_fallback_solution, :316-343
        const double e0 = x0x - r0x, e2 = b2c_wrap(x0t - r0t);
        u00 = b2c_clip(ur00 + (-1.0 * e0), -a.v_max, a.v_max);
        u01 = b2c_clip(ur01 + (-0.5 * e2), -a.w_max, a.w_max);
        Jout = (double)INFINITY;   // +inf
        max_viol = 0.0;
        for (int k = 0; k <= N; ++k) {
            if (a.xpred) {
                a.xpred[(int64_t)(3 * k) * a.B + i] = x0x;
                a.xpred[(int64_t)(3 * k + 1) * a.B + i] = x0y;
                a.xpred[(int64_t)(3 * k + 2) * a.B + i] = x0t;
            }
            if (k < N && a.useq) {
                a.useq[(int64_t)(2 * k) * a.B + i] = u00;
                a.useq[(int64_t)(2 * k + 1) * a.B + i] = u01;
            }
        }
    }
    a.u0[i] = u00; a.u0[a.B + i] = u01;
    a.cost[i] = Jout;
    if (a.status) a.status[i] = st;
    if (a.iters) a.iters[i] = it;
    if (a.slack_used) a.slack_used[i] = (max_viol > 1e-6) ? 1 : 0;   // :485 / :294-296
}

template <bool BS1>
__global__ void __launch_bounds__(MPC_TPB)
k_mpc(const MpcArgs a) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < mpc_count(a)) mpc_solve_one<BS1>(a, i);
}

}  // namespace

static int mpc_fields(const b2c_mpc_params* p) {
    const int bs = (p->kind == B2C_MPC_LTV && p->block_size > 1) ? p->block_size : 1;
    const int Nb = (p->N + bs - 1) / bs;
    const int lam = (!p->soft && p->M > 0) ? p->N * p->M : 0;      // one multiplier per hard row
    return (bs == 1 ? Lay<true>::total(p->N, Nb) : Lay<false>::total(p->N, Nb)) + lam;
}

int64_t b2c_mpc_scratch_bytes(const b2c_mpc_params* p, int64_t B) {
    if (!p || B <= 0) return 0;
    MpcArgs a;
    mpc_fill_args(a, B, p, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr,
                  nullptr, nullptr);
    int tpb = 0;
    if (p->engine != B2C_MPC_ENGINE_GENERAL && mpc_slots_eligible(a, &tpb)) {
        int dev = 0, sms = 148;                       // the device the calling thread is on (informational figure)
        if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        return mpc_slots_scratch_bytes(a, sms);
    }
    return (int64_t)mpc_fields(p) * B * (int64_t)sizeof(double);
}

static int mpc_check_params(const b2c_mpc_params* p) {
    if (p->N < 1 || p->N > 4096 || p->M < 0) { b2c_set_error("bad horizon / obstacle count"); return -2; }
    if (p->kind != B2C_MPC_LTV && p->kind != B2C_MPC_LTI) { b2c_set_error("bad MPC kind"); return -2; }
    if (!(p->Rd[0] > 0) || !(p->Rd[1] > 0)) { b2c_set_error("R must be positive definite"); return -3; }
    return 0;
}

bool mpc_uses_slots(const MpcArgs& a, int engine) {
    int tpb = 0;
    return engine != B2C_MPC_ENGINE_GENERAL && mpc_slots_eligible(a, &tpb);
}

int mpc_launch(b2c_ctx* ctx, MpcArgs& a, int engine) {
    int tpb = 0;
    const bool slots_ok = mpc_slots_eligible(a, &tpb);
    if (engine == B2C_MPC_ENGINE_SLOTS && !slots_ok) {
        b2c_set_error("slot kernel needs kind=LTV, block_size=1, soft rows and a horizon whose strips fit one SM");
        return -5;
    }
    if (slots_ok && engine != B2C_MPC_ENGINE_GENERAL) return mpc_slots_launch(ctx, a, tpb);
    if (a.in_aos || a.out_aos || a.table) { b2c_set_error("internal: the general kernel takes SoA buffers"); return -6; }
    b2c_mpc_params pp;
    memset(&pp, 0, sizeof pp);
    pp.kind = a.kind; pp.N = a.N; pp.M = a.M; pp.block_size = a.bs; pp.soft = a.soft;
    const size_t need = (size_t)mpc_fields(&pp) * (size_t)a.B * sizeof(double);
    int rc = b2c_reserve(&ctx->scratch, &ctx->scratch_bytes, need, false);
    if (rc) return rc;
    a.scratch = (double*)ctx->scratch;
    a.lam_off = mpc_fields(&pp) - ((!a.soft && a.M > 0) ? a.N * a.M : 0);
    if (a.bs == 1) k_mpc<true><<<b2c_blocks(a.B, MPC_TPB), MPC_TPB, 0, ctx->stream>>>(a);
    else           k_mpc<false><<<b2c_blocks(a.B, MPC_TPB), MPC_TPB, 0, ctx->stream>>>(a);
    B2C_CHECK_LAUNCH(ctx);
    return 0;
}

int b2c_mpc_dev(b2c_ctx* ctx, int64_t B, const b2c_mpc_params* p,
                const double* x0, const double* xref, const double* uref, const double* obs,
                int32_t* step_count,
                double* u0, double* cost, int32_t* status, int32_t* iters, int32_t* slack_used,
                double* useq, double* xpred) {
    if (!ctx || !p) { b2c_set_error("null context/params"); return -1; }
    int rc = mpc_check_params(p);
    if (rc) return rc;
    if (B <= 0) return 0;
    B2C_DEVICE_GUARD(ctx);
    MpcArgs a;
    mpc_fill_args(a, B, p, x0, xref, uref, obs, step_count, u0, cost, status, iters, slack_used, useq, xpred);
    return mpc_launch(ctx, a, p->engine);
}

// (x_refs, u_refs) of instance i = rows min(k0[i] + k, rows - 1) of the shared table -> instance-minor SoA (only the
// general kernel needs them materialised; the slot path gathers inside k_mpc_prepare)
__global__ void __launch_bounds__(256)
k_mpc_table_gather(const MpcArgs a, double* __restrict__ xref, double* __restrict__ uref) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, B = a.B;
    if (i >= mpc_count(a)) return;
    for (int k = 0; k <= a.N; ++k) {
        const double* row = a.table + (int64_t)mpc_table_row(a, i, k) * 6;
        xref[(int64_t)(3 * k) * B + i] = row[1]; xref[(int64_t)(3 * k + 1) * B + i] = row[2];
        xref[(int64_t)(3 * k + 2) * B + i] = row[3];
        if (k < a.N) { uref[(int64_t)(2 * k) * B + i] = row[4]; uref[(int64_t)(2 * k + 1) * B + i] = row[5]; }
    }
}

// table form with SoA x0 / obs / outputs (`a.table`, `a.table_rows`, `a.k0` set by the caller)
int mpc_table_launch(b2c_ctx* ctx, MpcArgs& a, int engine) {
    if (mpc_uses_slots(a, engine)) return mpc_launch(ctx, a, engine);
    const size_t nx = (size_t)3 * (a.N + 1) * a.B, nu = (size_t)2 * a.N * a.B;
    int rc = b2c_reserve(&ctx->aux, &ctx->aux_bytes, (nx + nu) * sizeof(double), false);
    if (rc) return rc;
    double* xr = (double*)ctx->aux; double* ur = xr + nx;
    k_mpc_table_gather<<<b2c_blocks(a.B, 256), 256, 0, ctx->stream>>>(a, xr, ur);
    B2C_CHECK_LAUNCH(ctx);
    a.xref = xr; a.uref = ur; a.table = nullptr; a.k0 = nullptr; a.table_rows = 0;
    return mpc_launch(ctx, a, engine);
}

int b2c_mpc_table_dev(b2c_ctx* ctx, int64_t B, const b2c_mpc_params* p,
                      const double* table, int32_t table_rows, const int32_t* k0,
                      const double* x0, const double* obs, int32_t* step_count,
                      double* u0, double* cost, int32_t* status, int32_t* iters, int32_t* slack_used,
                      double* useq, double* xpred) {
    if (!ctx || !p || !table || !k0) { b2c_set_error("null context/params/table"); return -1; }
    int rc = mpc_check_params(p);
    if (rc) return rc;
    if (table_rows < 1) { b2c_set_error("empty reference table"); return -2; }
    if (p->kind != B2C_MPC_LTV) { b2c_set_error("the table form is solve_with_ltv's (kind = B2C_MPC_LTV)"); return -2; }
    if (B <= 0) return 0;
    B2C_DEVICE_GUARD(ctx);
    MpcArgs a;
    mpc_fill_args(a, B, p, x0, nullptr, nullptr, obs, step_count, u0, cost, status, iters, slack_used, useq, xpred);
    a.table = table; a.table_rows = table_rows; a.k0 = k0;
    return mpc_table_launch(ctx, a, p->engine);
}

// ---- allocation-free variant of b2c_mpc_dev (no counterpart in the reference)
// ------------------------------------------------------------------------------------------------------
// For callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it): the
// general kernel k_mpc with its per-instance scratch in a file-scope device array instead of the context's arena.
// Every GPU that loads libb2c.so holds the array (16 MiB) until the process exits.  All calls on a device share it:
// calls that overlap on the GPU (different streams / contexts) race.  Nothing in it has to start at zero -- the
// kernel writes every field it later reads (linearization, inputs, gains, hard-row multipliers: mpc_solve_one's
// setup) -- so the contents a previous call leaves behind are never observed.
#define B2C_STATIC_SCRATCH_DOUBLES (2 * 1024 * 1024)
static __device__ __align__(256) double g_mpc_static_scratch[B2C_STATIC_SCRATCH_DOUBLES];

#ifndef B2C_HOST_EMUL
#include <mutex>
// In host code g_mpc_static_scratch is only a stand-in for the symbol: its device address is looked up once per
// device, with that device current, and THAT pointer goes to the kernel.
static int mpc_static_scratch_ptr(int device, double** out) {
    static std::mutex mu;
    static double* ptr[64] = {};
    if (device < 0 || device >= 64) { b2c_set_error("static scratch: device index out of range"); return -11; }
    std::lock_guard<std::mutex> lk(mu);
    if (!ptr[device]) {
        double* q = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&q, g_mpc_static_scratch);
        if (e != cudaSuccess) return b2c_cuda_fail(e, "cudaGetSymbolAddress(g_mpc_static_scratch)", __FILE__, __LINE__);
        ptr[device] = q;
    }
    *out = ptr[device];
    return 0;
}
#endif

int64_t b2c_mpc_static_max_instances(const b2c_mpc_params* p) {
    if (!p || p->N < 1 || p->M < 0) return 0;
    return (int64_t)B2C_STATIC_SCRATCH_DOUBLES / mpc_fields(p);
}

int b2c_mpc_dev_static(b2c_ctx* ctx, int64_t B, const b2c_mpc_params* p,
                       const double* x0, const double* xref, const double* uref, const double* obs,
                       int32_t* step_count,
                       double* u0, double* cost, int32_t* status, int32_t* iters, int32_t* slack_used,
                       double* useq, double* xpred) {
#ifndef B2C_HOST_EMUL
    if (!ctx || !p) { b2c_set_error("null context/params"); return -1; }
    if (p->N < 1 || p->N > 4096 || p->M < 0) { b2c_set_error("bad horizon / obstacle count"); return -2; }
    if (p->kind != B2C_MPC_LTV && p->kind != B2C_MPC_LTI) { b2c_set_error("bad MPC kind"); return -2; }
    if (!(p->Rd[0] > 0) || !(p->Rd[1] > 0)) { b2c_set_error("R must be positive definite"); return -3; }
    if (B <= 0) return 0;
    const int64_t fields = mpc_fields(p);
    if (fields * B > (int64_t)B2C_STATIC_SCRATCH_DOUBLES) {
        char buf[256];
        snprintf(buf, sizeof buf, "b2c_mpc_dev_static: %lld instances x %lld scratch doubles exceed the static array "
                 "(%d doubles = 16 MiB): at most %lld instances of this shape per call; use b2c_mpc_dev for more",
                 (long long)B, (long long)fields, B2C_STATIC_SCRATCH_DOUBLES,
                 (long long)((int64_t)B2C_STATIC_SCRATCH_DOUBLES / fields));
        b2c_set_error(buf);
        return -7;
    }
    B2C_DEVICE_GUARD(ctx);
    double* scratch = nullptr;
    int rc = mpc_static_scratch_ptr(ctx->device, &scratch);
    if (rc) return rc;
    MpcArgs a;
    mpc_fill_args(a, B, p, x0, xref, uref, obs, step_count, u0, cost, status, iters, slack_used, useq, xpred);
    a.scratch = scratch;
    a.lam_off = (int)fields - ((!p->soft && p->M > 0) ? p->N * p->M : 0);
    if (a.bs == 1) k_mpc<true><<<b2c_blocks(B, MPC_TPB), MPC_TPB, 0, ctx->stream>>>(a);
    else           k_mpc<false><<<b2c_blocks(B, MPC_TPB), MPC_TPB, 0, ctx->stream>>>(a);
    B2C_CHECK_LAUNCH(ctx);
    return 0;
#else
    (void)ctx; (void)B; (void)p; (void)x0; (void)xref; (void)uref; (void)obs; (void)step_count; (void)u0; (void)cost;
    (void)status; (void)iters; (void)slack_used; (void)useq; (void)xpred;
    return -1;
#endif
}

#ifdef B2C_HOST_EMUL
#include <vector>
// Test-only: the same per-instance solver on the host, SoA host pointers (see b2c_mpc_dev).
extern "C" int b2c_emul_mpc(int64_t B, const b2c_mpc_params* p,
                            const double* x0, const double* xref, const double* uref, const double* obs,
                            int32_t* step_count, double* u0, double* cost, int32_t* status,
                            int32_t* iters, int32_t* slack_used, double* useq, double* xpred) {
    MpcArgs a;
    mpc_fill_args(a, B, p, x0, xref, uref, obs, step_count, u0, cost, status, iters, slack_used, useq, xpred);
    std::vector<double> scratch((size_t)mpc_fields(p) * (size_t)B);
    a.scratch = scratch.data();
    a.lam_off = mpc_fields(p) - ((!p->soft && p->M > 0) ? p->N * p->M : 0);
    for (int64_t i = 0; i < B; ++i) {
        if (a.bs == 1) mpc_solve_one<true>(a, i); else mpc_solve_one<false>(a, i);
    }
    return 0;
}
#endif
