"""GPU parity: the CUDA path (through the C ABI / ctypes, libb2c.so) against the CPU oracle and the
golden fixtures.  Tolerances are BASELINE.json's: DARE gains <= 1e-9 relative; MPC first control
<= 1e-4, objective <= 1e-6 relative, constraint violation <= 1e-6; linearization <= 4e-16."""
import numpy as np
import pytest

from conftest import MPC_PRM6, MPC_PRM20, DEFAULT_OBS, g_mpc_open_loop_tol, assert_g_mpc_closed_loop

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def hc():
    import hybrid_controller
    from hybrid_controller import _native
    _native.load_library()
    return hybrid_controller


# ------------------------------------------------------------------------------------------ K1
def test_linearizer_matches_reference_golden(hc, golden):
    from hybrid_controller.models.linearization import Linearizer
    r = golden["ref_py"]
    lin = Linearizer(dt=float(r["dt"]))
    v, th = r["v"], r["th"]
    for meth, key in (("euler", "euler"), ("exact", "exact"), ("jacobian", "jac")):
        A, B = lin.get_discrete_model_batch(v, th, meth)
        assert np.abs(A - r[key + "_A"]).max() <= 4e-16, meth
        assert np.abs(B - r[key + "_B"]).max() <= 4e-16, meth
    A, B = lin.get_discrete_model_explicit_batch(v, th)
    assert np.abs(A - r["explicit_A"]).max() <= 4e-16 and np.abs(B - r["explicit_B"]).max() <= 4e-16
    # scalar API, shapes, error behaviour
    A1, B1 = lin.get_discrete_model_explicit(v[3], th[3])
    assert A1.shape == (3, 3) and B1.shape == (3, 2) and A1.dtype == np.float64
    assert np.array_equal(A1, A[3]) and np.array_equal(B1, B[3])
    Aj, Bj = lin.get_jacobians(v[3], th[3])
    assert np.abs(Aj - r["jac_A"][3]).max() <= 4e-16
    with pytest.raises(ValueError):
        lin.get_discrete_model(1.0, 0.0, method="tustin")
    # rollouts and condensed matrices
    tr = lin.predict_trajectory(r["roll_x0"], r["roll_ctrl"], r["roll_v"], r["roll_th"])
    assert np.abs(tr - r["roll_ltv"]).max() < 1e-14
    tr = lin.predict_horizon(r["roll_x0"], r["roll_ctrl"], r["roll_v"][0], r["roll_th"][0])
    assert np.abs(tr - r["roll_lti"]).max() < 1e-14
    Phi, Gam = lin.build_prediction_matrices(r["explicit_A"][5], r["explicit_B"][5], 4)
    assert np.abs(Phi - r["pm_Phi"]).max() < 1e-15 and np.abs(Gam - r["pm_Gamma"]).max() < 1e-15


def test_linearize_large_batch_and_ragged(hc):
    from hybrid_controller.models.linearization import Linearizer
    from oracle import lin_dare
    lin = Linearizer(dt=0.02)
    rng = np.random.default_rng(1)
    for B in (1, 31, 257, (1 << 20) + 3):
        v, th = rng.uniform(-2, 2, B), rng.uniform(-np.pi, np.pi, B)
        A, Bm = lin.get_discrete_model_explicit_batch(v, th)
        s, c = np.sin(th), np.cos(th)
        assert np.abs(A[:, 0, 2] - (-v * s * 0.02)).max() <= 4e-16
        assert np.abs(A[:, 1, 2] - (v * c * 0.02)).max() <= 4e-16
        assert np.abs(Bm[:, 0, 0] - c * 0.02).max() <= 4e-16 and np.abs(Bm[:, 1, 0] - s * 0.02).max() <= 4e-16
        assert np.all(A[:, 0, 0] == 1) and np.all(A[:, 2, 2] == 1) and np.all(A[:, 1, 0] == 0)
        assert np.all(Bm[:, 2, 1] == 0.02) and np.all(Bm[:, 0, 1] == 0)
    Ae, Be = lin_dare.discrete_explicit(v[0], th[0], 0.02)
    assert np.abs(A[0] - Ae).max() <= 4e-16


# ------------------------------------------------------------------------------------------ K2
@pytest.mark.parametrize("tag", ["sim", "def", "aniso"])
def test_dare_matches_reference_golden(hc, golden, tag):
    from hybrid_controller.controllers.lqr_controller import LQRController
    r = golden["ref_py"]
    lqr = LQRController(Q_diag=list(r[f"Qd_{tag}"]), R_diag=list(r[f"Rd_{tag}"]), dt=0.02, v_max=2.0,
                        omega_max=3.0)
    K, P, st = lqr.compute_gain_batch(r["v"], r["th"], want_P=True)
    assert (st == 0).all()
    Kref, Pref = r[f"K_{tag}"], r[f"P_{tag}"]
    relK = np.abs(K - Kref).max(axis=(1, 2)) / np.abs(Kref).max(axis=(1, 2))
    relP = np.abs(P - Pref).max(axis=(1, 2)) / np.abs(Pref).max(axis=(1, 2))
    assert relK.max() <= 1e-9, relK.max()
    assert relP.max() <= 1e-9, relP.max()


def test_dare_config3_vs_scipy_and_properties(hc):
    from hybrid_controller.controllers.lqr_controller import LQRController
    from hybrid_controller import workloads as wl
    from oracle import lin_dare
    lqr = LQRController(**wl.SIM_LQR)
    B = 1 << 20
    for stress in (False, True):
        v, th = wl.config3_points(B, stress=stress)
        K, P, st = lqr.compute_gain_batch(v, th, want_P=True)
        assert (st == 0).all()
        idx = np.linspace(0, B - 1, 257).astype(int)
        for i in idx:
            K0, P0 = lin_dare.lqr_gain(v[i], th[i], [15, 15, 8], [.1, .1], .02)
            assert np.abs(K[i] - K0).max() <= 1e-9 * np.abs(K0).max()
            assert np.abs(P[i] - P0).max() <= 1e-9 * np.abs(P0).max()
        # size-independent properties on the FULL batch: DARE residual and rotation equivariance
        c, s = np.cos(th), np.sin(th)
        A = np.tile(np.eye(3), (B, 1, 1)); A[:, 0, 2] = -v * s * .02; A[:, 1, 2] = v * c * .02
        Bm = np.zeros((B, 3, 2)); Bm[:, 0, 0] = c * .02; Bm[:, 1, 0] = s * .02; Bm[:, 2, 1] = .02
        Q, R = np.diag([15., 15., 8.]), np.diag([.1, .1])
        AtP = np.swapaxes(A, 1, 2) @ P
        res = AtP @ A - AtP @ Bm @ K + Q - P            # P = A'PA - A'PB K + Q
        assert (np.abs(res).max(axis=(1, 2)) / np.abs(P).max(axis=(1, 2))).max() < 1e-11
        assert np.abs(P - np.swapaxes(P, 1, 2)).max() == 0
        # q_x == q_y  =>  K(v,th) = K(v,0) Rot(-th): |K[:, :2]| row norms do not depend on th
        K0b, _ = lqr.compute_gain_batch(v[:4096], np.zeros(4096))
        n_rot = np.linalg.norm(K[:4096, :, :2], axis=2)
        n_0 = np.linalg.norm(K0b[:, :, :2], axis=2)
        assert np.abs(n_rot - n_0).max() <= 1e-9 * n_0.max()


def test_lqr_scalar_api_cache_guard_fallback(hc, capsys):
    from hybrid_controller.controllers.lqr_controller import LQRController
    from oracle import lin_dare
    lqr = LQRController(Q_diag=[15.0, 15.0, 8.0], R_diag=[0.1, 0.1], dt=0.02, v_max=2.0, omega_max=3.0)
    assert not lqr.gain_computed
    K = lqr.compute_gain(1.0, 0.0)
    assert np.allclose(K, [[10.838962679253070, 0, 0], [0, 11.058616855573257, 9.456836750988225]],
                       rtol=1e-10, atol=1e-12)
    assert lqr.gain_computed and lqr.P.shape == (3, 3)
    assert lqr.compute_gain(1.0 + 1e-7, 1e-7) is K                 # cache hit (:112-114)
    assert lqr.compute_gain(1.0 + 1e-7, 1e-7, force_recompute=True) is not K
    Kg = lqr.compute_gain(0.0, 0.3)                                # |v|<1e-6 guard -> v = 0.01
    K0, _ = lin_dare.lqr_gain(0.0, 0.3, [15, 15, 8], [.1, .1], .02)
    assert np.abs(Kg - K0).max() <= 1e-9 * np.abs(K0).max()
    with pytest.raises(np.linalg.LinAlgError):                     # get_lqr_gain: no guard (:232-237)
        lqr.get_lqr_gain(0.0, 0.3)
    K2 = lqr.get_lqr_gain(0.8, -1.0, dt=0.05)
    K20, _ = lin_dare.lqr_gain(0.8, -1.0, [15, 15, 8], [.1, .1], .05, guard=False)
    assert np.abs(K2 - K20).max() <= 1e-9 * np.abs(K20).max()
    lqr.set_weights(Q_diag=[1.0, 1.0, 1.0])
    assert lqr.K is None and lqr.P is None
    # NaN operating point -> reference prints a warning and uses [[1,0,0],[0,0,1]] (:134-141)
    Kf = lqr.compute_gain(float("nan"), 0.0, force_recompute=True)
    assert np.array_equal(Kf, [[1.0, 0, 0], [0, 0, 1.0]])
    assert "DARE solver failed" in capsys.readouterr().out


def test_lqr_control_law_and_g_lqr_closed_loop(hc, golden):
    """Config 1: run_lqr_simulation's loop with the drop-in classes vs the shipped log G-LQR."""
    from hybrid_controller.controllers.lqr_controller import LQRController
    from hybrid_controller.models.differential_drive import DifferentialDriveRobot
    from hybrid_controller.trajectory.reference_generator import ReferenceTrajectoryGenerator
    r, g = golden["ref_py"], golden["g_lqr"]
    lqr = LQRController(Q_diag=[15.0, 15.0, 8.0], R_diag=[0.1, 0.1], dt=0.02, v_max=2.0, omega_max=3.0)
    ub, eb = lqr.compute_control_batch(r["law_x"], r["law_xr"], r["law_ur"])
    assert np.abs(ub - r["law_u"]).max() < 1e-8 and np.abs(eb - r["law_e"]).max() < 1e-15
    robot = DifferentialDriveRobot(v_max=2.0, omega_max=3.0)
    gen = ReferenceTrajectoryGenerator(A=2.0, a=0.5, dt=0.02)
    gen.generate(20.0)
    x = gen.get_reference_at_index(0)[0].copy()
    worst_x = worst_u = 0.0
    for k in range(999):
        x_ref, u_ref = gen.get_reference_at_index(k)
        u, e = lqr.compute_control_at_operating_point(x, x_ref, u_ref)
        worst_x = max(worst_x, np.abs(x - g["states"][k]).max())
        worst_u = max(worst_u, np.abs(u - g["controls"][k]).max())
        x = robot.simulate_step(x, u, 0.02)
    assert worst_x <= 1e-9 and worst_u <= 1e-9, (worst_x, worst_u)


# ------------------------------------------------------------------------------------------ K3
def _check_against_oracle(out, b, o, tol_u=1e-4, tol_j=1e-6):
    assert out["status"][b] == 0
    assert np.abs(out["u0"][b] - o["optimal_control"]).max() <= tol_u
    assert abs(out["cost"][b] - o["cost"]) <= tol_j * abs(o["cost"])
    assert bool(out["slack_used"][b]) == bool(o["slack_used"])
    assert np.abs(out["useq"][b] - o["control_sequence"]).max() <= 1e-4
    assert np.abs(out["xpred"][b] - o["predicted_states"]).max() <= 1e-5


def _violations(mpc, x0, xr, ur, ob, out, b, bs=1):
    """max residual of dynamics / input box / (soft) half-space rows at the returned point."""
    from oracle import mpc_qp
    N = mpc.N
    prm = dict(N=N, Qd=np.diag(mpc.Q), Rd=np.diag(mpc.R), Pd=np.diag(mpc.P), d_safe=mpc.d_safe,
               rho=mpc.slack_penalty, v_max=mpc.v_max, w_max=mpc.omega_max, dt=mpc.dt, block_size=bs)
    qp = mpc_qp.build_ltv(x0[b], xr[b], ur[b], ob[b], **prm)
    Nb = qp.meta["Nb"]
    xu = xr[b].copy(); xu[:, 2] = np.unwrap(xu[:, 2])
    # dx from xpred needs the reference's branch: recompute it through the equality rows instead
    du = out["useq"][b] - ur[b][:N]
    dub = du[::bs][:Nb]
    z = np.zeros(qp.H.shape[0])
    z[qp.meta["u_off"]:qp.meta["u_off"] + 2 * Nb] = dub.ravel()
    nxs = 3 * (N + 1)
    z[:nxs] = np.linalg.solve(qp.Aeq[:, :nxs], qp.beq - qp.Aeq[:, nxs:] @ z[nxs:])
    # predicted states agree with the dynamics roll-out (mod the wrapped/unwrapped reference branch)
    dxk = out["xpred"][b] - xr[b][:N + 1]
    assert np.abs(dxk[:, :2] - z[:nxs].reshape(N + 1, 3)[:, :2]).max() <= 1e-6
    box = max((np.abs(out["useq"][b][:, 0]) - mpc.v_max).max(), (np.abs(out["useq"][b][:, 1]) - mpc.omega_max).max())
    rows = qp.G[4 * N:] @ z - qp.h[4 * N:]           # obstacle rows with slack = 0
    slack = np.maximum(0.0, -rows[:, ]) if len(rows) else np.zeros(0)
    J = 0.5 * z @ qp.H @ z + qp.c @ z + qp.c0 + mpc.slack_penalty * np.sum(slack ** 2)
    return max(box, 0.0), J


def _one_blas_thread():
    # the pool already uses every core: without this each worker's BLAS spawns a thread per core (800 s instead of 5)
    try:
        from threadpoolctl import threadpool_limits
        _one_blas_thread.limit = threadpool_limits(1)
    except Exception:
        pass


def _oracle_config4(args):
    from oracle import mpc_qp
    x0, xr, ur, ob = args
    o = mpc_qp.solve_with_ltv(x0, xr, ur, ob, **MPC_PRM20)
    return {k: o[k] for k in ("optimal_control", "cost", "slack_used", "control_sequence", "predicted_states")}


@pytest.fixture(scope="module")
def config4_oracle():
    """the 1,024-instance config-4 subsample, every instance solved by the oracle (all host cores;
    the pool is forked before this module touches CUDA)."""
    import multiprocessing as mp
    import os
    from hybrid_controller import workloads as wl
    B = 1024
    x0, xr, ur, ob = wl.config4_instances(B, N=20, seed=0)
    cores = len(os.sched_getaffinity(0))
    with mp.get_context("fork").Pool(cores, initializer=_one_blas_thread) as pool:
        res = pool.map(_oracle_config4, [(x0[b], xr[b], ur[b], ob[b]) for b in range(B)], chunksize=8)
    return (x0, xr, ur, ob), res


@pytest.mark.parametrize("engine", [1, 2])
def test_mpc_config4_vs_oracle(config4_oracle, hc, engine):
    """Config 4 shape (N=20, M=5, block 1): ALL 1,024 instances of the subsample against the exact oracle,
    through both K3 kernels (1 = general HBM-scratch kernel, 2 = persistent slot kernel)."""
    from hybrid_controller.controllers.mpc_controller import MPCController
    from hybrid_controller import workloads as wl
    (x0, xr, ur, ob), oracle = config4_oracle
    B = len(oracle)
    mpc = MPCController(horizon=20, block_size=1, **wl.SIM_MPC)
    mpc.engine = engine
    out = mpc.solve_with_ltv_batch(x0, xr, ur, ob)
    assert (out["status"] == 0).all()
    assert out["slack_used"].sum() > 0 and (np.abs(out["u0"][:, 0]) >= 2.0 - 1e-12).sum() > 0
    for b in range(B):
        _check_against_oracle(out, b, oracle[b])
        if b % 8 == 0:
            box, J = _violations(mpc, x0, xr, ur, ob, out, b)
            assert box <= 1e-6 and abs(J - out["cost"][b]) <= 1e-6 * abs(J)


@pytest.mark.parametrize("N,bs,M", [(6, 2, 3), (6, 1, 3), (20, 3, 5), (7, 4, 1), (10, 1, 0), (1, 1, 2)])
def test_mpc_shapes_blocking_vs_oracle(hc, N, bs, M):
    from hybrid_controller.controllers.mpc_controller import MPCController
    from hybrid_controller import workloads as wl
    from oracle import mpc_qp
    B = 96
    mpc = MPCController(horizon=N, block_size=bs, **wl.SIM_MPC)
    x0, xr, ur, ob = wl.config4_instances(B, N=N, seed=N * 10 + bs, obstacles=wl.DENSE_OBSTACLES[:M],
                                          pos_jitter=0.2, th_jitter=0.5)
    out = mpc.solve_with_ltv_batch(x0, xr, ur, ob)
    prm = dict(MPC_PRM20, N=N, block_size=bs)
    for b in range(B):
        o = mpc_qp.solve_with_ltv(x0[b], xr[b], ur[b], ob[b], **prm)
        _check_against_oracle(out, b, o)


def test_mpc_g_mpc_golden_log(hc, golden):
    """Config 2: the 200 solves of the shipped run log G-MPC through the scalar drop-in API."""
    from hybrid_controller.controllers.mpc_controller import MPCController, Obstacle
    from hybrid_controller.trajectory.reference_generator import ReferenceTrajectoryGenerator
    g = golden["g_mpc"]
    mpc = MPCController(horizon=6, Q_diag=[15.0, 15.0, 50.0], R_diag=[0.1, 0.1], P_diag=[30.0, 30.0, 40.0],
                        d_safe=0.3, slack_penalty=5000.0, dt=0.02, v_max=2.0, omega_max=3.0,
                        solver='OSQP', block_size=2)
    obstacles = [Obstacle(x=1.0, y=0.5, radius=0.2), Obstacle(x=-0.5, y=-1.0, radius=0.25),
                 Obstacle(x=1.5, y=-0.3, radius=0.15)]
    gen = ReferenceTrajectoryGenerator(A=2.0, a=0.5, dt=0.02)
    gen.generate(20.0)
    slack_ticks = set(g["slack_ticks"].tolist())
    seen = []
    for k in range(0, 999, 5):
        x_refs, u_refs = gen.get_trajectory_segment(k, mpc.N + 1)
        sol = mpc.solve_with_ltv(g["states"][k].copy(), x_refs, u_refs, obstacles)
        assert sol.status == "optimal" and sol.iterations == 0
        assert sol.control_sequence.shape == (6, 2) and sol.predicted_states.shape == (7, 3)
        err = np.abs(sol.optimal_control - g["controls"][k]).max()
        assert err <= g_mpc_open_loop_tol(k, slack_ticks), (k, err)   # per-tick bands, two named OSQP outliers
        if sol.slack_used:
            seen.append(k)
    assert seen == sorted(slack_ticks)
    assert mpc._step_count == 200
    ws = mpc.get_warm_start()
    assert ws.shape == (6, 2) and np.array_equal(ws[:-1], mpc._prev_solution[1:])
    mpc.reset()
    assert mpc._step_count == 0 and mpc.get_warm_start() is None


def test_mpc_lti_solve_vs_oracle(hc):
    from hybrid_controller.controllers.mpc_controller import MPCController, Obstacle
    from hybrid_controller import workloads as wl
    from oracle import mpc_qp
    mpc = MPCController(horizon=10, Q_diag=[10.0, 10.0, 1.0], R_diag=[0.1, 0.1], P_diag=[20.0, 20.0, 2.0],
                        d_safe=0.3, slack_penalty=1000.0, v_max=1.0, omega_max=1.5, dt=0.02, solver="ECOS")
    prm = dict(N=10, Qd=[10, 10, 1], Rd=[.1, .1], Pd=[20, 20, 2], d_safe=.3, rho=1000., v_max=1., w_max=1.5, dt=.02)
    table = wl.figure8_table()
    obstacles = [Obstacle(*o) for o in DEFAULT_OBS]
    rng = np.random.default_rng(5)
    for k0, nrows in ((0, 11), (120, 11), (400, 5), (995, 11), (45, 11), (52, 3)):
        idx = np.minimum(k0 + np.arange(nrows), len(table) - 1)
        xr, ur = table[idx, 1:4], table[idx, 4:6][:max(nrows - 1, 1)]
        x0 = xr[0] + rng.uniform(-.1, .1, 3)
        sol = mpc.solve(x0, xr, ur, obstacles)
        o = mpc_qp.solve(x0, xr, ur, DEFAULT_OBS, **prm)
        assert sol.status == o["status"] == "optimal"
        assert np.abs(sol.optimal_control - o["optimal_control"]).max() <= 1e-4
        assert abs(sol.cost - o["cost"]) <= 1e-6 * abs(o["cost"])
        assert sol.slack_used == o["slack_used"]
        assert np.abs(sol.predicted_states - o["predicted_states"]).max() <= 1e-5


def test_mpc_ramp_fallback_and_edge_cases(hc):
    from hybrid_controller.controllers.mpc_controller import MPCController, Obstacle
    from hybrid_controller import workloads as wl
    from oracle import mpc_qp
    mpc = MPCController(horizon=6, block_size=2, **wl.SIM_MPC)
    x0, xr, ur, ob = wl.config4_instances(8, N=6, seed=3, obstacles=wl.DEFAULT_OBSTACLES, th_jitter=1.0)
    # cold-start ramp (mpc_controller.py:500-507) with per-instance counters
    sc = np.arange(8, dtype=np.int32) * 2
    out = mpc.solve_with_ltv_batch(x0, xr, ur, ob, step_count=sc)
    assert np.array_equal(out["step_count"], sc + 1)
    for b in range(8):
        o = mpc_qp.solve_with_ltv(x0[b], xr[b], ur[b], ob[b], step_count=int(sc[b]), **MPC_PRM6)
        assert np.abs(out["u0"][b] - o["optimal_control"]).max() <= 1e-4
        assert np.abs(out["useq"][b] - o["control_sequence"]).max() <= 1e-4
    # NaN state -> fallback semantics (:316-343): status fallback, cost inf, tiled outputs
    xb = x0.copy(); xb[2, 0] = np.nan
    out = mpc.solve_with_ltv_batch(xb, xr, ur, ob, step_count=sc)
    assert out["status"][2] != 0 and np.isinf(out["cost"][2]) and out["step_count"][2] == sc[2]
    assert (out["status"][[0, 1, 3, 4, 5, 6, 7]] == 0).all()
    sol = mpc.solve_with_ltv(np.array([0.3, 0.1, 7.0]) * np.array([1, 1, np.inf]), xr[0], ur[0], [])
    assert sol.status == "fallback" and sol.cost == float("inf") and not sol.slack_used
    fb = mpc_qp.fallback([0.3, 0.1, 0.5], xr[0], ur[0], 6, 2.0, 3.0)
    sol2 = mpc._get_fallback_solution(np.array([0.3, 0.1, 0.5]), xr[0], ur[0], 0.0)
    assert np.allclose(sol2.optimal_control, fb["optimal_control"])
    # theta reference crossing +-pi inside the horizon and x0 on another 2pi branch (:392-401)
    table = wl.figure8_table()
    k0 = int(np.argmax(np.abs(np.diff(table[:, 3])) > 3.0)) - 2
    xr2, ur2 = table[k0:k0 + 7, 1:4], table[k0:k0 + 6, 4:6]
    x02 = xr2[0] + [0.02, -0.01, 2 * np.pi + 0.05]
    sol = mpc.solve_with_ltv(x02, xr2, ur2, [Obstacle(*o) for o in DEFAULT_OBS])
    o = mpc_qp.solve_with_ltv(x02, xr2, ur2, DEFAULT_OBS, step_count=0, **MPC_PRM6)
    assert np.abs(sol.optimal_control - o["optimal_control"]).max() <= 1e-4
    assert abs(sol.cost - o["cost"]) <= 1e-6 * abs(o["cost"])
    # obstacle centre on a reference point: row skipped (dist <= 0.01, Q7); v_ref guard (Q2)
    ob3 = [Obstacle(xr2[2, 0], xr2[2, 1] + 0.005, 0.2)]
    ur3 = ur2.copy(); ur3[1, 0] = 0.005; ur3[3, 0] = -0.004
    sol = mpc.solve_with_ltv(xr2[0].copy(), xr2, ur3, ob3)
    o = mpc_qp.solve_with_ltv(xr2[0], xr2, ur3, [(ob3[0].x, ob3[0].y, 0.2)], step_count=1, **MPC_PRM6)
    assert np.abs(sol.optimal_control - o["optimal_control"]).max() <= 1e-4
    assert abs(sol.cost - o["cost"]) <= 1e-6 * abs(o["cost"])
    with pytest.raises(IndexError):
        mpc.solve_with_ltv(xr2[0], xr2[:4], ur2[:3], [])


@pytest.mark.parametrize("N,M,B", [(20, 5, 20000), (6, 3, 300), (10, 0, 77), (1, 2, 33), (32, 3, 4000), (40, 5, 5000), (85, 2, 700)])
def test_mpc_engines_agree(hc, N, M, B):
    """The slot kernel (shared-memory scratch, lane refill; 128 / 64 / 32 slots per CTA depending on N)
    and the general kernel run the same method: same iteration counts, results equal to rounding.
    Includes ragged batch sizes, per-instance ramp counters and a NaN instance (fallback law)."""
    from hybrid_controller.controllers.mpc_controller import MPCController
    from hybrid_controller import workloads as wl
    x0, xr, ur, ob = wl.config4_instances(B, N=N, seed=N + M, obstacles=wl.DENSE_OBSTACLES[:M],
                                          pos_jitter=0.15, th_jitter=0.4)
    x0[B // 2, 1] = np.nan
    sc = (np.arange(B) % 13).astype(np.int32)
    outs = []
    for engine in (1, 2):
        mpc = MPCController(horizon=N, block_size=1, **wl.SIM_MPC)
        mpc.engine = engine
        outs.append(mpc.solve_with_ltv_batch(x0, xr, ur, ob, step_count=sc.copy()))
    a, b = outs
    assert np.array_equal(a["status"], b["status"]) and a["status"][B // 2] != 0 and (a["status"] != 0).sum() == 1
    assert np.array_equal(a["step_count"], b["step_count"]) and np.array_equal(a["slack_used"], b["slack_used"])
    assert np.abs(a["iters"] - b["iters"]).max() <= 1
    ok = a["status"] == 0
    assert np.abs(a["u0"] - b["u0"]).max() <= 1e-9 and np.abs(a["useq"] - b["useq"]).max() <= 1e-9
    assert np.abs(a["xpred"][ok] - b["xpred"][ok]).max() <= 1e-9
    assert np.array_equal(a["xpred"][~ok], b["xpred"][~ok], equal_nan=True)
    assert (np.abs(a["cost"][ok] - b["cost"][ok]) <= 1e-10 * np.abs(a["cost"][ok])).all()
    assert np.isinf(b["cost"][~ok]).all()


def test_mpc_full_size_properties(hc):
    """65,536 instances (BASELINE config 4 size): size-independent checks on the whole batch."""
    from hybrid_controller.controllers.mpc_controller import MPCController
    from hybrid_controller import workloads as wl
    B = 65536
    mpc = MPCController(horizon=20, block_size=1, **wl.SIM_MPC)
    x0, xr, ur, ob = wl.config4_instances(B, N=20, seed=0)
    out = mpc.solve_with_ltv_batch(x0, xr, ur, ob)
    assert (out["status"] == 0).all()
    u = out["useq"]
    assert np.abs(u[:, :, 0]).max() <= 2.0 + 1e-12 and np.abs(u[:, :, 1]).max() <= 3.0 + 1e-12
    assert np.array_equal(u[:, 0, :], out["u0"])
    # idempotence / determinism: a second launch is bit-identical
    out2 = mpc.solve_with_ltv_batch(x0, xr, ur, ob)
    assert np.array_equal(out2["u0"], out["u0"]) and np.array_equal(out2["cost"], out["cost"])
    # sharding invariance (8e): a slice solved alone equals the same global indices of the full batch
    s, c = wl.shard_range(B, 3, 8)
    xs = wl.config4_instances(B, N=20, seed=0, start=s, count=c)
    outs = mpc.solve_with_ltv_batch(*xs)
    assert np.array_equal(outs["u0"], out["u0"][s:s + c]) and np.array_equal(outs["cost"], out["cost"][s:s + c])
    # the no-obstacle, unsaturated subset is a pure LQ problem: 1 Newton step
    assert out["iters"].min() >= 1 and out["iters"].max() <= 30


def test_mpc_config2_closed_loop_vs_g_mpc(hc, golden):
    """Config 2: run_mpc_simulation's loop with the drop-in classes (batch = 1 latency path) vs G-MPC.
    Bands as in the CPU twin (conftest.assert_g_mpc_closed_loop): 1e-9 up to tick 29, 5e-4 / 5e-3 after the log's
    unpolished k=30 solve has entered its own loop."""
    import time
    from hybrid_controller.controllers.mpc_controller import MPCController, Obstacle
    from hybrid_controller.models.differential_drive import DifferentialDriveRobot
    from hybrid_controller.trajectory.reference_generator import ReferenceTrajectoryGenerator
    g = golden["g_mpc"]
    robot = DifferentialDriveRobot(v_max=2.0, omega_max=3.0)
    gen = ReferenceTrajectoryGenerator(A=2.0, a=0.5, dt=0.02)
    gen.generate(20.0)
    mpc = MPCController(horizon=6, Q_diag=[15.0, 15.0, 50.0], R_diag=[0.1, 0.1], P_diag=[30.0, 30.0, 40.0],
                        d_safe=0.3, slack_penalty=5000.0, dt=0.02, v_max=2.0, omega_max=3.0, solver='OSQP',
                        block_size=2)
    obstacles = [Obstacle(x=1.0, y=0.5, radius=0.2), Obstacle(x=-0.5, y=-1.0, radius=0.25),
                 Obstacle(x=1.5, y=-0.3, radius=0.15)]
    x = gen.get_reference_at_index(0)[0].copy()
    xs, us = [], []
    slack, lat = [], []
    for k in range(999):
        x_refs, u_refs = gen.get_trajectory_segment(k, mpc.N + 1)
        if k % 5 == 0:
            t0 = time.perf_counter()
            sol = mpc.solve_with_ltv(x, x_refs, u_refs, obstacles)
            lat.append((time.perf_counter() - t0) * 1e3)
            assert sol.status == "optimal"
            if sol.slack_used:
                slack.append(k)
        u = sol.optimal_control
        xs.append(np.array(x, float)); us.append(np.array(u, float))
        x = robot.simulate_step(x, u, 0.02)
    assert_g_mpc_closed_loop(xs, us, g["states"], g["controls"])
    assert slack == g["slack_ticks"].tolist()
    lat = np.array(lat[5:])
    print(f"\nbatch=1 solve_with_ltv latency: mean {lat.mean():.3f} ms  p50 {np.median(lat):.3f}  p99 "
          f"{np.percentile(lat, 99):.3f} ms   (reference log: mean 82.6 ms, p99 154.5 ms)")
    assert np.median(lat) < 5.0


@pytest.mark.gpu
@pytest.mark.parametrize("N,bs,M", [(20, 1, 5), (6, 2, 3)])
def test_mpc_hard_rows_vs_oracle_and_emulation(hc, N, bs, M):
    """use_soft_constraints=False on the device (method of multipliers in k_mpc): batch against the host emulation of
    the same code, a subset against the oracle's exact QP, and the scalar drop-in call incl. an infeasible instance
    (status "fallback", cost inf, as the reference returns when its solver reports infeasible)."""
    import emul
    from oracle import mpc_qp
    from hybrid_controller.controllers.mpc_controller import MPCController, Obstacle
    from hybrid_controller import workloads as wl, _native as nat
    B = 2048
    x0, xr, ur, ob = wl.config4_instances(B, N=N, seed=77, obstacles=wl.DENSE_OBSTACLES[:M])
    mpc = MPCController(horizon=N, block_size=bs, **wl.SIM_MPC)
    out = mpc.solve_with_ltv_batch(x0, xr, ur, ob, use_soft_constraints=False)
    p = nat.mpc_params(nat.MPC_LTV, N, M, bs, [15, 15, 50], [.1, .1], [30, 30, 40], .3, 5000., 2., 3., .02, soft=False)
    ref = emul.mpc(p, x0, xr, ur, ob)
    assert np.array_equal(out["status"], ref["status"]) and set(np.unique(out["status"])) == {0, 3}
    ok = out["status"] == 0
    assert np.abs(out["u0"] - ref["u0"]).max() <= 1e-7 and np.abs(out["xpred"][ok] - ref["xpred"][ok]).max() <= 1e-7
    assert (np.abs(out["cost"][ok] - ref["cost"][ok]) <= 1e-8 * np.abs(ref["cost"][ok])).all()
    assert not out["slack_used"].any() and np.isinf(out["cost"][~ok]).all()
    prm = dict(MPC_PRM20, N=N, block_size=bs, soft=False)
    seen_fb = False
    for b in list(range(24)) + np.nonzero(~ok)[0][:3].tolist():
        o = mpc_qp.solve_with_ltv(x0[b], xr[b], ur[b], ob[b], **prm)
        assert (o["status"] == "optimal") == bool(ok[b])
        assert np.abs(out["u0"][b] - o["optimal_control"]).max() <= 1e-4
        if ok[b]:
            assert abs(out["cost"][b] - o["cost"]) <= 1e-6 * abs(o["cost"])
        else:
            seen_fb = True
            sol = mpc.solve_with_ltv(x0[b], xr[b], ur[b], [Obstacle(*r) for r in ob[b]], use_soft_constraints=False)
            assert sol.status == "fallback" and np.isinf(sol.cost) and not sol.slack_used
            assert np.allclose(sol.optimal_control, o["optimal_control"])
    assert seen_fb


@pytest.mark.gpu
def test_mpc_lti_hard_rows_scalar_api(hc):
    """MPCController.solve (LTI, absolute state, padded references) with use_soft_constraints=False through the
    drop-in class, against the oracle's exact QP; both verdicts occur."""
    from oracle import mpc_qp
    from hybrid_controller.controllers.mpc_controller import MPCController, Obstacle
    from hybrid_controller import workloads as wl
    N, M = 8, 3
    x0, xr, ur, ob = wl.config4_instances(40, N=N, seed=9, obstacles=wl.DEFAULT_OBSTACLES)
    mpc = MPCController(horizon=N, **wl.SIM_MPC)
    prm = dict(MPC_PRM20, N=N, block_size=1, soft=False)
    verdicts = set()
    for b in range(40):
        obs = [Obstacle(*r) for r in ob[b]]
        sol = mpc.solve(x0[b], xr[b][:5], ur[b][:4], obs, use_soft_constraints=False)      # short refs: padded, :172-183
        o = mpc_qp.solve(x0[b], xr[b][:5], ur[b][:4], ob[b], **prm)
        assert sol.status == o["status"] and not sol.slack_used
        verdicts.add(sol.status)
        assert np.abs(sol.optimal_control - o["optimal_control"]).max() <= 1e-4
        if sol.status == "optimal":
            assert abs(sol.cost - o["cost"]) <= 1e-6 * abs(o["cost"])
            assert np.abs(sol.predicted_states - o["predicted_states"]).max() <= 1e-5
    assert verdicts == {"optimal", "fallback"}


@pytest.mark.gpu
@pytest.mark.parametrize("N,M,B", [(20, 5, 3000), (6, 3, 700), (20, 5, 40)])
def test_mpc_slot_kernel_helper_lanes_follow_the_sequential_line_search(hc, N, M, B):
    """Helper lanes (mpc_slots.cu: idle lanes of a warp evaluate the next trial steps of an instance whose trial was
    rejected) must reproduce the SEQUENTIAL line search: same iteration count and status per instance as the host
    emulation of the per-lane state machine (which has no helpers), results equal to rounding of the device's
    sincos / reciprocal.  The batch sits deep inside the obstacles (large start offsets), so most instances see
    rejected trials, and it is small enough for the whole launch to run in tail mode (helpers reserved from the
    first round on)."""
    import emul
    from hybrid_controller.controllers.mpc_controller import MPCController
    from hybrid_controller import workloads as wl, _native as nat
    x0, xr, ur, ob = wl.config4_instances(B, N=N, seed=5 * N + M, obstacles=wl.DENSE_OBSTACLES[:M],
                                          pos_jitter=0.35, th_jitter=0.6, obs_jitter=0.3)
    mpc = MPCController(horizon=N, block_size=1, **wl.SIM_MPC)
    mpc.engine = 2
    out = mpc.solve_with_ltv_batch(x0, xr, ur, ob)
    p = nat.mpc_params(nat.MPC_LTV, N, M, 1, [15, 15, 50], [.1, .1], [30, 30, 40], .3, 5000., 2., 3., .02)
    ref = emul.mpc(p, x0, xr, ur, ob, slots=True)
    assert np.array_equal(out["status"], ref["status"]) and (out["status"] == 0).all()
    same = out["iters"] == ref["iters"]
    # an iteration count may differ only where the decrement sits at the stopping threshold (device vs host libm)
    assert same.mean() >= 0.995 and np.abs(out["iters"] - ref["iters"]).max() <= 1
    assert out["iters"].max() >= 5                               # the batch does contain long line searches
    assert np.abs(out["u0"] - ref["u0"]).max() <= 1e-9
    assert (np.abs(out["cost"] - ref["cost"]) <= 1e-10 * np.abs(ref["cost"])).all()
    assert np.abs(out["useq"] - ref["useq"]).max() <= 1e-8 and np.abs(out["xpred"] - ref["xpred"]).max() <= 1e-8
    again = mpc.solve_with_ltv_batch(x0, xr, ur, ob)             # helpers or not, a relaunch reproduces every bit
    for key in ("u0", "cost", "iters", "useq", "xpred", "slack_used"):
        assert np.array_equal(out[key], again[key]), key


# ------------------------------------------------------------------------------------------ round 2: table form, async, slot configs
def _table_batch(B, N, M, seed):
    from hybrid_controller import workloads as wl
    table = wl.figure8_table()
    rng = np.random.default_rng(seed)
    k0 = rng.integers(0, 1000, B).astype(np.int32)
    k0[:8] = [999, 998, 990, 985, 980, 979, 0, 1]                     # segments clamped at the last table row
    idx = np.minimum(k0[:, None] + np.arange(N + 1)[None], len(table) - 1)
    xr, ur = table[idx][:, :, 1:4].copy(), table[idx][:, :N, 4:6].copy()
    x0 = xr[:, 0] + rng.uniform(-1, 1, (B, 3)) * [.1, .1, .2]
    ob = np.broadcast_to(wl.DENSE_OBSTACLES[:M][None], (B, M, 3)).copy()
    ob[:, :, :2] += rng.uniform(-.2, .2, (B, M, 2))
    return table, k0, x0, xr, ur, ob


@pytest.mark.parametrize("N,M,bs,B", [(20, 5, 1, 70000), (6, 3, 1, 3000), (6, 3, 2, 900), (20, 5, 1, 33)])
def test_mpc_table_form_equals_materialised_segments(hc, N, M, bs, B):
    """b2c_mpc_table_host (segments gathered on the device from the shared table, get_trajectory_segment's clamp,
    HC/trajectory/reference_generator.py:299-326) == b2c_mpc_host on the materialised segments, every bit -- through
    the slot path (chunked for the large batch) and, with move blocking, through the general kernel."""
    from hybrid_controller.controllers.mpc_controller import MPCController
    from hybrid_controller import workloads as wl
    table, k0, x0, xr, ur, ob = _table_batch(B, N, M, seed=N + B)
    x0[B // 2, 0] = np.nan
    sc = (np.arange(B) % 12).astype(np.int32)
    mpc = MPCController(horizon=N, block_size=bs, **wl.SIM_MPC)
    a = mpc.solve_with_ltv_batch(x0, xr, ur, ob, step_count=sc)
    b = mpc.solve_with_ltv_table_batch(x0, table, k0, ob, step_count=sc)
    assert (a["status"] != 0).sum() == 1 and a["status"][B // 2] != 0
    for k in ("u0", "cost", "status", "iters", "slack_used", "useq", "xpred", "step_count"):
        assert np.array_equal(a[k], b[k], equal_nan=True), k
    c = mpc.solve_with_ltv_table_batch(x0, table, k0, ob, step_count=sc, want_seq=False)
    assert np.array_equal(c["u0"], a["u0"], equal_nan=True) and c["useq"] is None


def test_mpc_host_async_two_calls_in_flight(hc):
    """b2c_mpc_host_async / b2c_mpc_table_host_async + b2c_wait: three calls issued back to back (the third waits for
    the first's staging arena) give the results of the synchronous calls."""
    import ctypes as C
    from hybrid_controller import _native as nat
    N, M, B = 20, 5, 20000
    ctx = nat.Context(0)
    prm = nat.mpc_params(nat.MPC_LTV, N, M, 1, [15, 15, 50], [.1, .1], [30, 30, 40], .3, 5000., 2., 3., .02)
    sets, outs = [], []
    for s in range(3):
        table, k0, x0, xr, ur, ob = _table_batch(B, N, M, seed=50 + s)
        sets.append((table, k0, x0, xr, ur, ob))
        outs.append(dict(u0=np.full((B, 2), np.nan), cost=np.full(B, np.nan), st=np.full(B, -1, np.int32),
                         it=np.full(B, -1, np.int32), sl=np.full(B, -1, np.int32), useq=np.full((B, N, 2), np.nan)))
    P = nat._ptr
    for s, ((table, k0, x0, xr, ur, ob), o) in enumerate(zip(sets, outs)):
        if s == 1:
            ctx.call("b2c_mpc_table_host_async", B, C.byref(prm), P(table), table.shape[0], P(k0), P(x0), P(ob), None,
                     P(o["u0"]), P(o["cost"]), P(o["st"]), P(o["it"]), P(o["sl"]), P(o["useq"]), None)
        else:
            ctx.call("b2c_mpc_host_async", B, C.byref(prm), P(x0), P(xr), P(ur), P(ob), None,
                     P(o["u0"]), P(o["cost"]), P(o["st"]), P(o["it"]), P(o["sl"]), P(o["useq"]), None)
    ctx.call("b2c_wait")
    for (table, k0, x0, xr, ur, ob), o in zip(sets, outs):
        ref = nat.mpc_host(ctx, prm, x0, xr, ur, ob)
        assert (o["st"] == 0).all()
        assert np.array_equal(o["u0"], ref["u0"]) and np.array_equal(o["cost"], ref["cost"])
        assert np.array_equal(o["useq"], ref["useq"]) and np.array_equal(o["it"], ref["iters"])
    ctx.close()


@pytest.mark.parametrize("N,M,B", [(20, 5, 50000), (20, 5, 3000), (6, 3, 9000)])
def test_mpc_slot_configurations_agree_bit_for_bit(hc, N, M, B, monkeypatch):
    """The slot kernel's configurations -- 256 slots per SM with the first 16 stages of the U strips in tensor memory
    (tcgen05.ld / tcgen05.st), 128 slots per SM with the strips in shared memory and the gains mirror in tensor or
    global memory, 256 in shared memory for short horizons -- run the same per-lane code on different storage:
    identical results, bit for bit."""
    from hybrid_controller.controllers.mpc_controller import MPCController
    from hybrid_controller import workloads as wl
    x0, xr, ur, ob = wl.config4_instances(B, N=N, seed=N + M + B, obstacles=wl.DENSE_OBSTACLES[:M],
                                          pos_jitter=0.25, th_jitter=0.5, obs_jitter=0.25)
    mpc = MPCController(horizon=N, block_size=1, **wl.SIM_MPC)
    mpc.engine = 2
    res = {}
    cfgs = ["256,16", "128,0", "128,32", "64,0"] + (["256,0"] if N <= 6 else [])
    for cfg in cfgs:
        monkeypatch.setenv("B2C_SLOTS_CFG", cfg)
        res[cfg] = mpc.solve_with_ltv_batch(x0, xr, ur, ob)
    # the gains mirror of the shared-memory configurations: tensor memory (default) against global memory
    monkeypatch.setenv("B2C_SLOTS_CFG", "128,0")
    monkeypatch.setenv("B2C_SLOTS_MIRROR", "0")
    res["128,0 / mirror in global memory"] = mpc.solve_with_ltv_batch(x0, xr, ur, ob)
    cfgs.append("128,0 / mirror in global memory")
    monkeypatch.delenv("B2C_SLOTS_MIRROR")
    monkeypatch.delenv("B2C_SLOTS_CFG")
    ref = res[cfgs[0]]
    assert (ref["status"] == 0).all() and ref["iters"].max() >= 5
    for cfg in cfgs[1:]:
        for k in ("u0", "cost", "iters", "slack_used", "useq", "xpred"):
            assert np.array_equal(ref[k], res[cfg][k]), (cfg, k)
