#!/usr/bin/env python3
"""Algorithm prototype (development aid, not product, not oracle): projected semismooth Newton
with a Riccati inner solve for the slack-eliminated MPC problem -- the method the CUDA kernel
implements.  Single instance, plain numpy; used to study iteration counts against oracle/mpc_qp."""
import numpy as np


def setup_ltv(x0, x_refs, u_refs, obstacles, *, N, Qd, Rd, Pd, d_safe, rho, v_max, w_max, dt,
              block_size=1, **_):
    xu = np.array(x_refs, float)
    xu[:, 2] = np.unwrap(xu[:, 2])
    d = x0[2] - xu[0, 2]
    while d > np.pi: d -= 2 * np.pi
    while d < -np.pi: d += 2 * np.pi
    z0 = np.array([x0[0] - xu[0, 0], x0[1] - xu[0, 1], d])
    A = np.zeros((N, 3, 3)); B = np.zeros((N, 3, 2))
    for k in range(N):
        v = u_refs[k, 0] if abs(u_refs[k, 0]) > 0.01 else 0.1
        s, c = np.sin(xu[k, 2]), np.cos(xu[k, 2])
        A[k] = [[1, 0, -v * s * dt], [0, 1, v * c * dt], [0, 0, 1]]
        B[k] = [[c * dt, 0], [s * dt, 0], [0, dt]]
    rows = [[] for _ in range(N)]
    for (ox, oy, r) in obstacles:
        for k in range(N):
            dx, dy = xu[k, 0] - ox, xu[k, 1] - oy
            dist = np.hypot(dx, dy)
            if dist > 0.01:
                rows[k].append((dx / dist, dy / dist, d_safe + r - dist))
    Nb = (N + block_size - 1) // block_size
    umax = np.array([v_max, w_max])
    lb = np.full((Nb, 2), -np.inf); ub = np.full((Nb, 2), np.inf)
    for k in range(N):
        j = min(k // block_size, Nb - 1)
        lb[j] = np.maximum(lb[j], -umax - u_refs[k]); ub[j] = np.minimum(ub[j], umax - u_refs[k])
    return dict(z0=z0, A=A, B=B, rows=rows, zc=np.zeros((N + 1, 3)), wc=-np.array(u_refs[:N], float),
                lb=lb, ub=ub, N=N, Nb=Nb, bs=block_size, Q=np.diag(Qd).astype(float),
                R=np.diag(Rd).astype(float), P=np.diag(Pd).astype(float), rho=rho)


def forward(p, w):
    """Rollout; returns cost (full, un-halved) and per-super-stage quadratic model (halved)."""
    N, Nb, bs = p["N"], p["Nb"], p["bs"]
    Q, R, P, rho = p["Q"], p["R"], p["P"], p["rho"]
    z = p["z0"].copy(); J = 0.0
    st = []
    nact = 0
    for j in range(Nb):
        Phi = np.eye(3); Gam = np.zeros((3, 2))
        Hzz = np.zeros((3, 3)); Hzu = np.zeros((3, 2)); Huu = np.zeros((2, 2)); gz = np.zeros(3); gu = np.zeros(2)
        for k in range(j * bs, min((j + 1) * bs, N)):
            e = z - p["zc"][k]
            W = Q.copy(); g = Q @ e
            J += e @ Q @ e
            for (nx, ny, h) in p["rows"][k]:
                viol = h - nx * z[0] - ny * z[1]
                if viol > 0:
                    nact += 1
                    J += rho * viol * viol
                    n = np.array([nx, ny, 0.0])
                    W += rho * np.outer(n, n); g -= rho * viol * n
            du = w[j] - p["wc"][k]
            J += du @ R @ du
            Hzz += Phi.T @ W @ Phi; Hzu += Phi.T @ W @ Gam; Huu += Gam.T @ W @ Gam + R
            gz += Phi.T @ g; gu += Gam.T @ g + R @ du
            z = p["A"][k] @ z + p["B"][k] @ w[j]
            Gam = p["A"][k] @ Gam + p["B"][k]; Phi = p["A"][k] @ Phi
        st.append((Hzz, Hzu, Huu, gz, gu, Phi, Gam))
    e = z - p["zc"][N]
    J += e @ P @ e
    return J, st, P @ e, nact


def solve(p, max_iter=50, verbose=False):
    Nb = p["Nb"]
    w = np.clip(np.zeros((Nb, 2)), p["lb"], p["ub"])
    J, st, lamN, _ = forward(p, w)
    for it in range(max_iter):
        S = p["P"].copy(); s = lamN.copy(); lam = lamN.copy()
        K = np.zeros((Nb, 2, 3)); kf = np.zeros((Nb, 2)); dec = 0.0
        for j in range(Nb - 1, -1, -1):
            Hzz, Hzu, Huu, gz, gu, Ab, Bb = st[j]
            g_true = gu + Bb.T @ lam
            lam = gz + Ab.T @ lam
            clamp = ((w[j] <= p["lb"][j]) & (g_true > 0)) | ((w[j] >= p["ub"][j]) & (g_true < 0))
            Qzz = Hzz + Ab.T @ S @ Ab; Quz = Hzu.T + Bb.T @ S @ Ab; Quu = Huu + Bb.T @ S @ Bb
            qz = gz + Ab.T @ s; qu = gu + Bb.T @ s
            for i in range(2):
                if clamp[i]:
                    Quu[i, :] = 0; Quu[:, i] = 0; Quu[i, i] = 1; Quz[i, :] = 0; qu[i] = 0
            K[j] = np.linalg.solve(Quu, Quz); kf[j] = np.linalg.solve(Quu, qu)
            S = Qzz - Quz.T @ K[j]; s = qz - Quz.T @ kf[j]
            S = 0.5 * (S + S.T)
            dec += 0.5 * qu @ kf[j]
        if verbose:
            print(it, J, dec)
        if dec <= 1e-14 * max(1.0, abs(J)):
            return w, J, it, True
        alpha = 1.0
        ok = False
        ntry = 0
        open_loop = False
        while ntry < 3 + 14:
            if ntry == 3 and not open_loop:
                # closed-loop clipped roll-outs failed: switch to the projection arc P[w + alpha d]
                # with d the Newton direction of the LINEAR model (Bertsekas 1982: guaranteed descent)
                dz = np.zeros(3); d = np.zeros_like(w)
                for j in range(Nb):
                    d[j] = -kf[j] - K[j] @ dz
                    dz = st[j][5] @ dz + st[j][6] @ d[j]
                open_loop = True; alpha = 1.0
            wn = np.zeros_like(w)
            z = p["z0"].copy(); zn = p["z0"].copy()
            for j in range(Nb):
                if open_loop:
                    wn[j] = np.clip(w[j] + alpha * d[j], p["lb"][j], p["ub"][j])
                else:
                    wn[j] = np.clip(w[j] - alpha * kf[j] - K[j] @ (zn - z), p["lb"][j], p["ub"][j])
                for k in range(j * p["bs"], min((j + 1) * p["bs"], p["N"])):
                    z = p["A"][k] @ z + p["B"][k] @ w[j]
                    zn = p["A"][k] @ zn + p["B"][k] @ wn[j]
            Jn, stn, lamNn, _ = forward(p, wn)
            if Jn < J:
                ok = True
                break
            alpha *= 0.5
            ntry += 1
        if not ok:
            return w, J, it, False
        w, J, st, lamN = wn, Jn, stn, lamNn
    return w, J, max_iter, False


def config4(n, seed=0, N=20):
    """SURVEY 8(d) config 4 instances (dense obstacles, jittered)."""
    import sys, os
    sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
    from oracle import lin_dare
    traj = lin_dare.figure8(20.0)
    rng = np.random.default_rng(seed)
    dense = np.array([[1.0, .5, .2], [-.5, -1.0, .25], [1.5, -.3, .15], [-1.5, .5, .2], [0, .8, .15]])
    out = []
    for _ in range(n):
        k0 = int(rng.integers(0, 980))
        xr, ur = lin_dare.segment(traj, k0, N + 1)
        x0 = xr[0] + np.array([rng.uniform(-.1, .1), rng.uniform(-.1, .1), rng.uniform(-.2, .2)])
        ob = dense.copy(); ob[:, :2] += rng.uniform(-.2, .2, (5, 2))
        out.append((x0, xr, ur, ob))
    return out


if __name__ == "__main__":
    import sys, os, time
    sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
    from oracle import mpc_qp
    prm = dict(N=20, Qd=[15, 15, 50], Rd=[.1, .1], Pd=[30, 30, 40], d_safe=.3, rho=5000., v_max=2.,
               w_max=3., dt=.02, block_size=int(sys.argv[2]) if len(sys.argv) > 2 else 1)
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 50
    its = []; worst = (0, 0)
    for i, (x0, xr, ur, ob) in enumerate(config4(n)):
        p = setup_ltv(x0, xr, ur, ob, **prm)
        w, J, it, ok = solve(p)
        o = mpc_qp.solve_with_ltv(x0, xr, ur, ob, **prm)
        u0 = ur[0] + w[0]
        e_u = np.abs(u0 - o["optimal_control"]).max(); e_J = abs(J - o["cost"]) / abs(o["cost"])
        its.append(it)
        if not ok or e_u > 1e-8 or e_J > 1e-10:
            print("inst", i, "ok", ok, "it", it, "e_u", e_u, "e_J", e_J, "J", J, o["cost"], o["slack_used"])
        worst = (max(worst[0], e_u), max(worst[1], e_J))
    print("iters: mean %.2f max %d hist %s" % (np.mean(its), max(its), np.bincount(its)), "worst", worst)
