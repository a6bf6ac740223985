"""Oracle: the MPC QP of the reference, restated densely and solved exactly.  TEST INFRASTRUCTURE.

Follows /root/reference/src/hybrid_controller/hybrid_controller/controllers/mpc_controller.py
  * ``build_ltv``  <- solve_with_ltv  :345-522 (LTV, error state, move blocking)
  * ``build_lti``  <- solve           :150-314 (LTI, absolute state, ref padding)
line by line into   min 1/2 z'Hz + c'z + c0   s.t.  Aeq z = beq,  G z >= h,  lo <= z <= hi
with the same variables the CVXPY model declares (states, [blocked] inputs, N*M slacks).

The reference hands this model to CVXPY -> OSQP (absent here, un-vendored dependency
``cvxpy>=1.4``, setup.py:24).  OSQP is an ADMM method stopped at eps=1e-5; a strictly convex
QP has ONE minimiser, so the oracle computes that minimiser to machine precision instead:
HiGHS' active-set QP solver (bundled in scipy) finds the active set, then the equality-
constrained KKT system of that active set is solved directly in fp64 ("polish") and its
multiplier signs / feasibility are verified.  Pinned against the shipped OSQP run log
(tests/golden/g_mpc.npz): see tests/test_oracle_golden.py.
"""
from dataclasses import dataclass
import numpy as np

from .lin_dare import discrete_explicit, wrap_angle

try:  # private scipy module: feature-probe, never assume
    import scipy.optimize._highspy._core as _hs
except Exception:  # pragma: no cover
    _hs = None


@dataclass
class QP:
    H: np.ndarray
    c: np.ndarray
    c0: float
    Aeq: np.ndarray
    beq: np.ndarray
    G: np.ndarray      # G z >= h
    h: np.ndarray
    lo: np.ndarray
    hi: np.ndarray
    meta: dict


def _obstacle_rows(N, nx_off, n_z, s_off, pref, obstacles, d_safe, soft, absolute):
    """mpc_controller.py:238-270 (absolute) / :439-468 (error form).

    Row: n.(p_k - p_obs) >= safe - s  with p_k = z_k (absolute) or pref_k + dz_k (error)."""
    rows, rhs = [], []
    sidx = 0
    for (ox, oy, orad) in obstacles:
        for k in range(N):
            dxo, dyo = pref[k, 0] - ox, pref[k, 1] - oy
            dist = np.sqrt(dxo ** 2 + dyo ** 2)
            if dist > 0.01:
                nx, ny = dxo / dist, dyo / dist
                safe = d_safe + orad
                r = np.zeros(n_z)
                r[nx_off + 3 * k + 0] = nx
                r[nx_off + 3 * k + 1] = ny
                if absolute:
                    b = safe + nx * ox + ny * oy
                else:
                    b = safe - (nx * (pref[k, 0] - ox) + ny * (pref[k, 1] - oy))
                if soft:
                    r[s_off + sidx] = 1.0
                    sidx += 1
                rows.append(r)
                rhs.append(b)
    return rows, rhs, sidx


def build_ltv(x0, x_refs, u_refs, obstacles, *, N, Qd, Rd, Pd, d_safe, rho, v_max, w_max, dt,
              block_size=1, soft=True):
    """solve_with_ltv, mpc_controller.py:345-480.  obstacles: iterable of (x, y, radius)."""
    x0 = np.asarray(x0, float)
    x_refs = np.asarray(x_refs, float)
    u_refs = np.asarray(u_refs, float)
    obstacles = [tuple(map(float, o)) for o in (obstacles if obstacles is not None else [])]
    M = len(obstacles)
    Nb = (N + block_size - 1) // block_size
    use_slack = soft and M > 0
    n_s = N * M if use_slack else 0
    nx_off, u_off, s_off = 0, 3 * (N + 1), 3 * (N + 1) + 2 * Nb
    n_z = s_off + n_s
    Q, R, P = np.diag(Qd).astype(float), np.diag(Rd).astype(float), np.diag(Pd).astype(float)

    xu = x_refs.copy()
    xu[:, 2] = np.unwrap(x_refs[:, 2])                       # :392-393
    th0 = xu[0, 2]
    x0a = x0.copy()
    x0a[2] = th0 + wrap_angle(x0[2] - th0)                   # :397-401

    H = np.zeros((n_z, n_z))
    c = np.zeros(n_z)
    c0 = 0.0
    blk = lambda k: min(k // block_size, Nb - 1)             # :374-380
    for k in range(N):                                       # :403-409
        H[3 * k:3 * k + 3, 3 * k:3 * k + 3] += 2 * Q
        j = u_off + 2 * blk(k)
        H[j:j + 2, j:j + 2] += 2 * R
        c[j:j + 2] += 2 * R @ u_refs[k]
        c0 += u_refs[k] @ R @ u_refs[k]
    H[3 * N:3 * N + 3, 3 * N:3 * N + 3] += 2 * P             # :412
    if n_s:
        H[s_off:, s_off:] += 2 * rho * np.eye(n_s)           # :415

    Aeq = np.zeros((3 * (N + 1), n_z))
    beq = np.zeros(3 * (N + 1))
    Aeq[:3, :3] = np.eye(3)
    beq[:3] = x0a - xu[0]                                    # :421
    for k in range(N):                                       # :424-428
        v_r = u_refs[k, 0] if abs(u_refs[k, 0]) > 0.01 else 0.1
        Ad, Bd = discrete_explicit(v_r, xu[k, 2], dt)
        r0 = 3 * (k + 1)
        Aeq[r0:r0 + 3, 3 * (k + 1):3 * (k + 1) + 3] = np.eye(3)
        Aeq[r0:r0 + 3, 3 * k:3 * k + 3] = -Ad
        j = u_off + 2 * blk(k)
        Aeq[r0:r0 + 3, j:j + 2] += -Bd

    rows, rhs = [], []
    umax = np.array([v_max, w_max])
    for k in range(N):                                       # :431-436
        j = u_off + 2 * blk(k)
        for i in range(2):
            r = np.zeros(n_z); r[j + i] = 1.0
            rows.append(r); rhs.append(-umax[i] - u_refs[k, i])
            rows.append(-r); rhs.append(-umax[i] + u_refs[k, i])
    orows, orhs, n_used = _obstacle_rows(N, nx_off, n_z, s_off, xu, obstacles, d_safe,
                                         use_slack, absolute=False)
    rows += orows; rhs += orhs
    lo = np.full(n_z, -np.inf); hi = np.full(n_z, np.inf)
    if n_s:
        lo[s_off:] = 0.0                                     # nonneg=True :384
    G = np.array(rows).reshape(-1, n_z)
    return QP(H, c, c0, Aeq, beq, G, np.array(rhs), lo, hi,
              dict(kind="ltv", N=N, Nb=Nb, M=M, bs=block_size, u_off=u_off, s_off=s_off,
                   n_s=n_s, x_refs=x_refs, u_refs=u_refs))


def build_lti(x0, x_refs, u_refs, obstacles, *, N, Qd, Rd, Pd, d_safe, rho, v_max, w_max, dt,
              soft=True, **_):
    """solve, mpc_controller.py:150-283."""
    x0 = np.asarray(x0, float)
    x_refs = np.asarray(x_refs, float)
    u_refs = np.asarray(u_refs, float)
    if x_refs.shape[0] < N + 1:                              # :172-177
        x_refs = np.vstack([x_refs, np.repeat(x_refs[-1:], N + 1 - x_refs.shape[0], 0)])
    if u_refs.shape[0] < N:                                  # :179-183
        u_refs = np.vstack([u_refs, np.repeat(u_refs[-1:], N - u_refs.shape[0], 0)])
    obstacles = [tuple(map(float, o)) for o in (obstacles if obstacles is not None else [])]
    M = len(obstacles)
    use_slack = soft and M > 0
    n_s = N * M if use_slack else 0
    u_off, s_off = 3 * (N + 1), 3 * (N + 1) + 2 * N
    n_z = s_off + n_s
    Q, R, P = np.diag(Qd).astype(float), np.diag(Rd).astype(float), np.diag(Pd).astype(float)
    v_r = u_refs[0, 0] if abs(u_refs[0, 0]) > 0.01 else 0.1  # :186
    Ad, Bd = discrete_explicit(v_r, x_refs[0, 2], dt)        # :187-190

    H = np.zeros((n_z, n_z)); c = np.zeros(n_z); c0 = 0.0
    for k in range(N):                                       # :206-210
        H[3 * k:3 * k + 3, 3 * k:3 * k + 3] += 2 * Q
        c[3 * k:3 * k + 3] += -2 * Q @ x_refs[k]
        c0 += x_refs[k] @ Q @ x_refs[k]
        j = u_off + 2 * k
        H[j:j + 2, j:j + 2] += 2 * R
    H[3 * N:3 * N + 3, 3 * N:3 * N + 3] += 2 * P             # :212-213
    c[3 * N:3 * N + 3] += -2 * P @ x_refs[N]
    c0 += x_refs[N] @ P @ x_refs[N]
    if n_s:
        H[s_off:, s_off:] += 2 * rho * np.eye(n_s)           # :216-217

    Aeq = np.zeros((3 * (N + 1), n_z)); beq = np.zeros(3 * (N + 1))
    Aeq[:3, :3] = np.eye(3); beq[:3] = x0                    # :223
    for k in range(N):                                       # :226-227
        r0 = 3 * (k + 1)
        Aeq[r0:r0 + 3, r0:r0 + 3] = np.eye(3)
        Aeq[r0:r0 + 3, 3 * k:3 * k + 3] = -Ad
        Aeq[r0:r0 + 3, u_off + 2 * k:u_off + 2 * k + 2] = -Bd
    rows, rhs = [], []
    umax = np.array([v_max, w_max])
    for k in range(N):                                       # :230-234
        for i in range(2):
            r = np.zeros(n_z); r[u_off + 2 * k + i] = 1.0
            rows.append(r); rhs.append(-umax[i])
            rows.append(-r); rhs.append(-umax[i])
    orows, orhs, _ = _obstacle_rows(N, 0, n_z, s_off, x_refs, obstacles, d_safe, use_slack,
                                    absolute=True)
    rows += orows; rhs += orhs
    lo = np.full(n_z, -np.inf); hi = np.full(n_z, np.inf)
    if n_s:
        lo[s_off:] = 0.0
    G = np.array(rows).reshape(-1, n_z)
    return QP(H, c, c0, Aeq, beq, G, np.array(rhs), lo, hi,
              dict(kind="lti", N=N, Nb=N, M=M, bs=1, u_off=u_off, s_off=s_off, n_s=n_s,
                   x_refs=x_refs, u_refs=u_refs))


# --------------------------------------------------------------------------- solvers
def _highs_solve(qp):
    """Active-set QP solve with HiGHS (fp64); returns z or None."""
    if _hs is None:
        return None
    n = qp.H.shape[0]
    Arows = np.vstack([qp.Aeq, qp.G])
    rl = np.concatenate([qp.beq, qp.h])
    ru = np.concatenate([qp.beq, np.full(len(qp.h), _hs.kHighsInf)])
    hh = _hs._Highs()
    hh.setOptionValue("output_flag", False)
    for opt in ("primal_feasibility_tolerance", "dual_feasibility_tolerance"):
        hh.setOptionValue(opt, 1e-10)
    m = _hs.HighsModel()
    lp = m.lp_
    lp.num_col_, lp.num_row_ = n, Arows.shape[0]
    lp.col_cost_ = qp.c.copy()
    lp.offset_ = float(qp.c0)
    lp.col_lower_ = np.where(np.isfinite(qp.lo), qp.lo, -_hs.kHighsInf)
    lp.col_upper_ = np.where(np.isfinite(qp.hi), qp.hi, _hs.kHighsInf)
    lp.row_lower_, lp.row_upper_ = rl, ru
    lp.a_matrix_.format_ = _hs.MatrixFormat.kRowwise
    nz = Arows != 0
    lp.a_matrix_.start_ = np.concatenate([[0], np.cumsum(nz.sum(1))]).astype(np.int32)
    lp.a_matrix_.index_ = np.nonzero(nz)[1].astype(np.int32)
    lp.a_matrix_.value_ = Arows[nz]
    Hl = np.tril(qp.H)
    hs = m.hessian_
    hs.dim_ = n
    hs.format_ = _hs.HessianFormat.kTriangular
    # column-wise lower triangle
    cols_start, idx, val = [0], [], []
    for j in range(n):
        r = np.nonzero(Hl[:, j])[0]
        idx.extend(r.tolist()); val.extend(Hl[r, j].tolist()); cols_start.append(len(idx))
    hs.start_ = np.array(cols_start, np.int32)
    hs.index_ = np.array(idx, np.int32)
    hs.value_ = np.array(val, float)
    if hh.passModel(m) != _hs.HighsStatus.kOk:
        return None
    hh.run()
    if hh.getModelStatus() != _hs.HighsModelStatus.kOptimal:
        return None
    return np.array(hh.getSolution().col_value)


def _ipm_solve(qp, iters=80):
    """Dense fp64 primal-dual interior point (Mehrotra), independent of HiGHS.

    Used when HiGHS is unavailable and as a second opinion in the tests."""
    n = qp.H.shape[0]
    fin = np.isfinite(qp.lo)
    Gb = np.eye(n)[fin]
    G = np.vstack([qp.G, Gb]); h = np.concatenate([qp.h, qp.lo[fin]])
    assert not np.isfinite(qp.hi).any()
    A, b = qp.Aeq, qp.beq
    m, p = len(h), len(b)
    z = np.zeros(n); y = np.zeros(p); s = np.ones(m); lam = np.ones(m)
    for _ in range(iters):
        rd = qp.H @ z + qp.c - A.T @ y - G.T @ lam
        rp = A @ z - b
        rg = G @ z - s - h
        mu = s @ lam / m
        if max(np.abs(rd).max(), np.abs(rp).max(), np.abs(rg).max()) < 1e-11 and mu < 1e-13:
            break
        W = lam / s
        K = np.block([[qp.H + G.T @ (W[:, None] * G), -A.T], [-A, np.zeros((p, p))]])

        def step(rc):
            # rc = target for s*lam complementarity residual
            r1 = -rd + G.T @ ((-rc - lam * rg) / s)   # eliminates ds, dlam
            sol = np.linalg.solve(K, np.concatenate([r1, rp]))
            dz, dy = sol[:n], sol[n:]
            ds = G @ dz + rg
            dl = (-rc - lam * ds) / s
            return dz, dy, ds, dl

        dz, dy, ds, dl = step(s * lam)
        def amax(v, dv):
            neg = dv < 0
            return min(1.0, (-v[neg] / dv[neg]).min()) if neg.any() else 1.0
        a_aff = min(amax(s, ds), amax(lam, dl))
        mu_aff = (s + a_aff * ds) @ (lam + a_aff * dl) / m
        sigma = (mu_aff / mu) ** 3
        dz, dy, ds, dl = step(s * lam + ds * dl - sigma * mu)
        a = 0.995 * min(amax(s, ds), amax(lam, dl))
        a = min(a, 1.0)
        z += a * dz; y += a * dy; s += a * ds; lam += a * dl
    return z


def _polish(qp, z, tol=1e-6):
    """Solve the KKT system of z's active set exactly; verify; else return z."""
    n = len(z)
    act_rows = np.nonzero(qp.G @ z - qp.h < tol)[0]
    act_lo = np.nonzero(np.isfinite(qp.lo) & (z - qp.lo < tol))[0]
    E = [qp.Aeq, qp.G[act_rows]]
    e = [qp.beq, qp.h[act_rows]]
    if len(act_lo):
        E.append(np.eye(n)[act_lo]); e.append(qp.lo[act_lo])
    E = np.vstack(E); e = np.concatenate(e)
    K = np.block([[qp.H, -E.T], [E, np.zeros((E.shape[0], E.shape[0]))]])
    rhs = np.concatenate([-qp.c, e])
    try:
        sol = np.linalg.lstsq(K, rhs, rcond=None)[0]
        for _ in range(2):                      # iterative refinement (rho=5e3 makes K stiff)
            sol = sol + np.linalg.lstsq(K, rhs - K @ sol, rcond=None)[0]
    except np.linalg.LinAlgError:
        return z, False
    zp, mult = sol[:n], sol[n:]
    scale = 1.0 + np.abs(mult).max(initial=0.0)
    ok = (np.abs(zp - z).max() < 1e-4
          and (qp.G @ zp - qp.h).min(initial=0.0) > -1e-8
          and (zp - qp.lo)[np.isfinite(qp.lo)].min(initial=0.0) > -1e-8
          and mult[len(qp.beq):].min(initial=0.0) > -1e-9 * scale
          and np.abs(qp.H @ zp + qp.c - E.T @ mult).max() < 1e-9 * scale)
    return (zp, True) if ok else (z, False)


def solve_qp(qp, method="highs", polish=True):
    """Returns (z, objective, info) or (None, inf, info) when no optimum is found."""
    z = _highs_solve(qp) if method == "highs" else None
    used = "highs"
    if z is None:
        try:
            z = _ipm_solve(qp); used = "ipm"
        except np.linalg.LinAlgError:
            return None, np.inf, dict(solver="none", polished=False)
    polished = False
    if polish:
        z, polished = _polish(qp, z)
    feas = max(np.abs(qp.Aeq @ z - qp.beq).max(), -(qp.G @ z - qp.h).min(initial=0.0),
               -(z - qp.lo)[np.isfinite(qp.lo)].min(initial=0.0))
    if not np.isfinite(z).all() or feas > 1e-6:
        return None, np.inf, dict(solver=used, polished=polished, feas=feas)
    obj = 0.5 * z @ qp.H @ z + qp.c @ z + qp.c0
    return z, float(obj), dict(solver=used, polished=polished, feas=float(feas))


# --------------------------------------------------------------------------- API-shaped wrappers
def fallback(x0, x_refs, u_refs, N, v_max, w_max):
    """_get_fallback_solution, mpc_controller.py:316-343."""
    e = np.asarray(x0, float) - np.asarray(x_refs[0], float)
    e[2] = wrap_angle(e[2])
    u = np.asarray(u_refs[0], float) - np.array([[1.0, 0, 0], [0, 0, 0.5]]) @ e
    u = np.array([np.clip(u[0], -v_max, v_max), np.clip(u[1], -w_max, w_max)])
    return dict(status="fallback", optimal_control=u, control_sequence=np.tile(u, (N, 1)),
                predicted_states=np.tile(np.asarray(x0, float), (N + 1, 1)), cost=float("inf"),
                slack_used=False)


def solve_with_ltv(x0, x_refs, u_refs, obstacles, step_count=10, ramp_up_steps=10, method="highs",
                   **prm):
    """solve_with_ltv incl. un-blocking and the cold-start ramp, mpc_controller.py:482-522.

    Returns a dict with MPCSolution's fields (+ 'slack', 'step_count' after the call)."""
    qp = build_ltv(x0, x_refs, u_refs, obstacles, **prm)
    z, obj, info = solve_qp(qp, method)
    N, Nb, bs = prm["N"], qp.meta["Nb"], qp.meta["bs"]
    if z is None:
        out = fallback(x0, x_refs, u_refs, N, prm["v_max"], prm["w_max"])
        out["step_count"] = step_count
        return out
    dxv = z[:3 * (N + 1)].reshape(N + 1, 3)
    dub = z[qp.meta["u_off"]:qp.meta["u_off"] + 2 * Nb].reshape(Nb, 2)
    du = np.array([dub[min(k // bs, Nb - 1)] for k in range(N)])
    slack = z[qp.meta["s_off"]:]
    x_pred = np.asarray(x_refs, float)[:N + 1] + dxv           # :497 (wrapped refs)
    u_pred = np.asarray(u_refs, float)[:N] + du                # :498
    if step_count < ramp_up_steps:                             # :502-505
        lim = prm["w_max"] * (step_count + 1) / ramp_up_steps
        u_pred[0, 1] = np.clip(u_pred[0, 1], -lim, lim)
    return dict(status="optimal", optimal_control=u_pred[0].copy(), control_sequence=u_pred,
                predicted_states=x_pred, cost=obj, slack_used=bool(np.any(slack > 1e-6)),
                slack=slack, step_count=step_count + 1, info=info)


def solve(x0, x_refs, u_refs, obstacles, method="highs", **prm):
    """solve (LTI), mpc_controller.py:285-314."""
    qp = build_lti(x0, x_refs, u_refs, obstacles, **prm)
    z, obj, info = solve_qp(qp, method)
    N = prm["N"]
    if z is None:
        return fallback(x0, qp.meta["x_refs"], qp.meta["u_refs"], N, prm["v_max"], prm["w_max"])
    xv = z[:3 * (N + 1)].reshape(N + 1, 3)
    uv = z[qp.meta["u_off"]:qp.meta["u_off"] + 2 * N].reshape(N, 2)
    slack = z[qp.meta["s_off"]:]
    return dict(status="optimal", optimal_control=uv[0].copy(), control_sequence=uv,
                predicted_states=xv, cost=obj, slack_used=bool(np.any(slack > 1e-6)), slack=slack,
                info=info)
