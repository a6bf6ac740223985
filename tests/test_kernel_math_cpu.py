"""CPU: (1) libb2c.so loads and exports every symbol include/b2c.h declares (no compute call);
(2) the per-instance kernel code (mpc_solve_one / dare_solve_one / b2c_linearize_point), compiled
for the host into a TEST-ONLY library (tests/emul.py), agrees with the oracle -- so the kernel
arithmetic is covered without a GPU; the -m gpu tests then check the real device path."""
import os
import re

import numpy as np
import pytest

import emul
from conftest import MPC_PRM6, MPC_PRM20, DEFAULT_OBS, ROOT, g_mpc_open_loop_tol, assert_g_mpc_closed_loop
from oracle import lin_dare, mpc_qp


def test_abi_library_exports_every_declared_symbol():
    from hybrid_controller import _native as nat
    lib = nat.load_library()
    hdr = open(os.path.join(ROOT, "include", "b2c.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(b2c_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 20
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in b2c.h but not exported"
    assert declared == set(nat.SIGNATURES), declared ^ set(nat.SIGNATURES)
    assert lib.b2c_version() == 200


def test_product_has_no_host_emulation_path():
    import subprocess
    from hybrid_controller import _native as nat
    syms = subprocess.run(["nm", "-D", nat.LIB_PATH], capture_output=True, text=True).stdout
    assert "b2c_emul" not in syms


def test_no_gpu_means_loud_failure():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from hybrid_controller import _native as nat
    from hybrid_controller.controllers import LQRController
    with pytest.raises(nat.B2CUnavailable):
        LQRController().compute_gain(1.0, 0.0)


def test_emul_linearize_and_dare_vs_reference_golden(golden):
    r = golden["ref_py"]
    for meth, key in ((1, "euler"), (2, "exact"), (0, "jac")):
        A, B = emul.linearize(r["v"], r["th"], 0.02, method=meth)
        assert np.abs(A - r[key + "_A"]).max() <= 4e-16 and np.abs(B - r[key + "_B"]).max() <= 4e-16
    for tag in ("sim", "def", "aniso"):
        K, P, it, st = emul.dare(r["v"], r["th"], 0.02, r[f"Qd_{tag}"], r[f"Rd_{tag}"])
        assert (st == 0).all() and it.max() <= 30
        relK = np.abs(K - r[f"K_{tag}"]).max(axis=(1, 2)) / np.abs(r[f"K_{tag}"]).max(axis=(1, 2))
        relP = np.abs(P - r[f"P_{tag}"]).max(axis=(1, 2)) / np.abs(r[f"P_{tag}"]).max(axis=(1, 2))
        assert relK.max() <= 1e-9 and relP.max() <= 1e-9
    # no guard + v = 0: no stabilizing solution -> non-zero status and the reference's fallback gain
    K, _, _, st = emul.dare([0.0], [0.3], 0.02, [15, 15, 8], [.1, .1], guard=0)
    assert st[0] != 0 and np.array_equal(K[0], [[1, 0, 0], [0, 0, 1]])


def _params(N, M, bs, kind=0, **kw):
    from hybrid_controller import _native as nat
    d = dict(Qd=[15, 15, 50], Rd=[.1, .1], Pd=[30, 30, 40], d_safe=.3, rho=5000., v_max=2., w_max=3., dt=.02)
    d.update(kw)
    return nat.mpc_params(kind, N, M, bs, d["Qd"], d["Rd"], d["Pd"], d["d_safe"], d["rho"], d["v_max"],
                          d["w_max"], d["dt"])


@pytest.mark.parametrize("N,bs,M", [(20, 1, 5), (6, 2, 3), (20, 3, 5), (7, 4, 1), (10, 1, 0), (1, 1, 2)])
def test_emul_mpc_vs_oracle(N, bs, M):
    from hybrid_controller import workloads as wl
    B = 64
    x0, xr, ur, ob = wl.config4_instances(B, N=N, seed=N * 10 + bs, obstacles=wl.DENSE_OBSTACLES[:M],
                                          pos_jitter=0.2, th_jitter=0.5)
    out = emul.mpc(_params(N, M, bs), x0, xr, ur, ob)
    prm = dict(MPC_PRM20, N=N, block_size=bs)
    assert (out["status"] == 0).all()
    for b in range(B):
        o = mpc_qp.solve_with_ltv(x0[b], xr[b], ur[b], ob[b], **prm)
        assert np.abs(out["u0"][b] - o["optimal_control"]).max() <= 1e-4
        assert abs(out["cost"][b] - o["cost"]) <= 1e-6 * abs(o["cost"])
        assert bool(out["slack_used"][b]) == o["slack_used"]
        assert np.abs(out["useq"][b] - o["control_sequence"]).max() <= 1e-4
        assert np.abs(out["xpred"][b] - o["predicted_states"]).max() <= 1e-5


@pytest.mark.parametrize("N,M", [(20, 5), (6, 3), (10, 0), (1, 2), (40, 5)])
def test_emul_slot_state_machine_vs_oracle_and_general(N, M):
    """The slot kernel's per-lane state machine (mpc_slots.cu: union gains/model storage, in-place input
    update, undo of rejected trials from the gain mirror) run on the host, against the oracle and
    against the general kernel's per-instance code."""
    from hybrid_controller import workloads as wl
    B = 96
    x0, xr, ur, ob = wl.config4_instances(B, N=N, seed=N * 7 + M, obstacles=wl.DENSE_OBSTACLES[:M],
                                          pos_jitter=0.2, th_jitter=0.5)
    x0[5, 0] = np.nan
    sc = (np.arange(B) % 12).astype(np.int32)
    out = emul.mpc(_params(N, M, 1), x0, xr, ur, ob, step_count=sc, slots=True)
    gen = emul.mpc(_params(N, M, 1), x0, xr, ur, ob, step_count=sc)
    assert np.array_equal(out["status"], gen["status"]) and out["status"][5] != 0 and (out["status"] != 0).sum() == 1
    assert np.array_equal(out["iters"], gen["iters"]) and np.array_equal(out["step_count"], gen["step_count"])
    # NaN input: the fallback law clips NaN to NaN, as np.clip does in the reference (mpc_controller.py:554-559)
    assert np.isinf(out["cost"][5]) and np.array_equal(out["u0"][5], gen["u0"][5], equal_nan=True) and np.isnan(out["u0"][5, 0])
    prm = dict(MPC_PRM20, N=N, block_size=1)
    for b in range(B):
        if b == 5:
            continue
        assert np.abs(out["u0"][b] - gen["u0"][b]).max() <= 1e-10
        o = mpc_qp.solve_with_ltv(x0[b], xr[b], ur[b], ob[b], step_count=int(sc[b]), **prm)
        assert np.abs(out["u0"][b] - o["optimal_control"]).max() <= 1e-4
        assert abs(out["cost"][b] - o["cost"]) <= 1e-6 * abs(o["cost"])
        assert bool(out["slack_used"][b]) == o["slack_used"]
        assert np.abs(out["useq"][b] - o["control_sequence"]).max() <= 1e-4
        assert np.abs(out["xpred"][b] - o["predicted_states"]).max() <= 1e-5


def test_emul_slot_rejected_trials_are_undone():
    """Instances whose line search rejects trials (incl. the projection-arc fallback) take the
    restore path of the slot kernel; the result must still equal the general kernel's."""
    from hybrid_controller import workloads as wl
    x0, xr, ur, ob = wl.config4_instances(4096, N=20, seed=0)
    a = emul.mpc(_params(20, 5, 1), x0, xr, ur, ob)
    b = emul.mpc(_params(20, 5, 1), x0, xr, ur, ob, slots=True)
    assert (b["status"] == 0).all() and np.array_equal(a["iters"], b["iters"]) and a["iters"].max() >= 8
    assert np.abs(a["u0"] - b["u0"]).max() <= 1e-10 and np.abs(a["useq"] - b["useq"]).max() <= 1e-10
    assert (np.abs(a["cost"] - b["cost"]) <= 1e-12 * np.abs(a["cost"])).all()


def test_emul_slot_layouts_and_table_form_are_bit_identical():
    """The slot path reads its inputs in three forms -- instance-minor SoA (b2c_mpc_dev), numpy instance-major
    (the *_host entry points hand their staging buffers over as they are) and shared table + start row
    (b2c_mpc_table_*: get_trajectory_segment on the device, HC/trajectory/reference_generator.py:299-326, clamped at
    the end of the table) -- and must produce the same bits from each, incl. start rows whose segment runs off the table."""
    from hybrid_controller import workloads as wl
    table = wl.figure8_table()
    rng = np.random.default_rng(3)
    B, N, M = 600, 20, 5
    k0 = rng.integers(0, 1000, B).astype(np.int32)
    k0[:8] = [999, 998, 990, 985, 980, 979, 0, 1]                     # segments clamped at the last row
    idx = np.minimum(k0[:, None] + np.arange(N + 1)[None], len(table) - 1)
    xr, ur = table[idx][:, :, 1:4].copy(), table[idx][:, :N, 4:6].copy()
    x0 = xr[:, 0] + rng.uniform(-1, 1, (B, 3)) * [.1, .1, .2]
    x0[5, 2] += 2 * np.pi                                              # another 2 pi branch
    x0[17, 1] = np.nan                                                 # fallback instance
    ob = np.broadcast_to(wl.DENSE_OBSTACLES[None], (B, M, 3)).copy()
    ob[:, :, :2] += rng.uniform(-.2, .2, (B, M, 2))
    sc = (np.arange(B) % 12).astype(np.int32)
    p = _params(N, M, 1)
    soa = emul.mpc(p, x0, xr, ur, ob, step_count=sc, slots=True)
    aos = emul.mpc_slots_ex(p, x0, ob, xref=xr, uref=ur, step_count=sc)
    tab = emul.mpc_slots_ex(p, x0, ob, table=table, k0=k0, step_count=sc)
    assert soa["status"][17] != 0 and (soa["status"] != 0).sum() == 1 and soa["slack_used"].sum() > 0
    for k in ("u0", "cost", "status", "iters", "slack_used", "useq", "xpred", "step_count"):
        assert np.array_equal(soa[k], aos[k], equal_nan=True), k
        assert np.array_equal(soa[k], tab[k], equal_nan=True), k


def test_emul_mpc_g_mpc_golden(golden):
    g = golden["g_mpc"]
    traj = lin_dare.figure8(20.0)
    ks = np.arange(0, 999, 5)
    x0 = g["states"][ks]
    xr = np.stack([lin_dare.segment(traj, k, 7)[0] for k in ks])
    ur = np.stack([lin_dare.segment(traj, k, 7)[1][:6] for k in ks])
    ob = np.broadcast_to(np.array(DEFAULT_OBS)[None], (len(ks), 3, 3))
    out = emul.mpc(_params(6, 3, 2), x0, xr, ur, ob, step_count=np.arange(len(ks), dtype=np.int32))
    assert (out["status"] == 0).all()
    slack_ticks = set(g["slack_ticks"].tolist())
    err = np.abs(out["u0"] - g["controls"][ks]).max(axis=1)
    for k, e in zip(ks, err):
        assert e <= g_mpc_open_loop_tol(int(k), slack_ticks), (k, e)
    assert sorted(ks[out["slack_used"] == 1].tolist()) == sorted(slack_ticks)


def test_emul_mpc_lti_and_fallback():
    from hybrid_controller import workloads as wl
    table = wl.figure8_table()
    rng = np.random.default_rng(5)
    prm = dict(N=10, Qd=[10, 10, 1], Rd=[.1, .1], Pd=[20, 20, 2], d_safe=.3, rho=1000., v_max=1., w_max=1.5, dt=.02)
    for k0 in (0, 45, 120, 400, 985):
        idx = np.minimum(k0 + np.arange(11), len(table) - 1)
        xr, ur = table[idx, 1:4], table[idx, 4:6][:10]
        x0 = xr[0] + rng.uniform(-.1, .1, 3)
        out = emul.mpc(_params(10, 3, 1, kind=1, **{k: v for k, v in prm.items() if k != 'N'}), x0[None], xr[None], ur[None], np.array(DEFAULT_OBS)[None])
        o = mpc_qp.solve(x0, xr, ur, DEFAULT_OBS, **prm)
        assert out["status"][0] == 0
        assert np.abs(out["u0"][0] - o["optimal_control"]).max() <= 1e-4
        assert abs(out["cost"][0] - o["cost"]) <= 1e-6 * abs(o["cost"])
        assert np.abs(out["xpred"][0] - o["predicted_states"]).max() <= 1e-5
    xb = x0.copy(); xb[1] = np.nan
    out = emul.mpc(_params(10, 3, 1, kind=1, **{k: v for k, v in prm.items() if k != 'N'}), xb[None], xr[None], ur[None], np.array(DEFAULT_OBS)[None])
    assert out["status"][0] != 0 and np.isinf(out["cost"][0])


def test_emul_mpc_infeasible_blocked_bounds():
    """block of 2 whose reference inputs differ by more than the box width -> no feasible du."""
    xr = np.zeros((5, 3)); ur = np.array([[1.9, 0.0], [-1.9, 0.0], [0.5, 0], [0.5, 0]])
    p = _params(4, 0, 2, v_max=1.0)
    out = emul.mpc(p, np.zeros((1, 3)), xr[None], ur[None], np.zeros((1, 0, 3)))
    assert out["status"][0] == 3 and np.isinf(out["cost"][0])
    fb = mpc_qp.fallback(np.zeros(3), xr, ur, 4, 1.0, 3.0)
    assert np.allclose(out["u0"][0], fb["optimal_control"])


def test_emul_mpc_closed_loop_vs_g_mpc(golden):
    """Config 2 closed loop (run_mpc_simulation: solve every 5th tick, hold u, Euler plant) vs the
    shipped log.  Bands of conftest.assert_g_mpc_closed_loop: 1e-9 up to tick 29, then states 5e-4 / controls 5e-3
    (the logged k=30 solve is an unpolished OSQP iterate, 1.37e-3 off, and re-enters the log's own loop); slack
    ticks identical."""
    g = golden["g_mpc"]
    traj = lin_dare.figure8(20.0)
    obs = np.array(DEFAULT_OBS)
    prm = _params(6, 3, 2)
    x = traj[0, 1:4].copy()
    sc = np.zeros(1, np.int32)
    xs, us = [], []
    slack = []
    for k in range(999):
        if k % 5 == 0:
            xr, ur = lin_dare.segment(traj, k, 7)
            o = emul.mpc(prm, x[None], xr[None], ur[None, :6], obs[None], step_count=sc)
            sc, u = o["step_count"], o["u0"][0]
            if o["slack_used"][0]:
                slack.append(k)
        xs.append(x.copy()); us.append(u.copy())
        x = lin_dare.plant_step(x, u, .02, 2., 3.)
    assert_g_mpc_closed_loop(xs, us, g["states"], g["controls"])
    assert slack == g["slack_ticks"].tolist()


@pytest.mark.parametrize("kind,N,bs,M", [(0, 20, 1, 5), (0, 6, 2, 3), (0, 10, 3, 5), (1, 8, 1, 3)])
def test_emul_mpc_hard_rows_vs_oracle(kind, N, bs, M):
    """use_soft_constraints=False (mpc_controller.py:197-200, :260-270, :383-386, :464-468): the obstacle rows are
    hard; an infeasible QP (e.g. the k = 0 row on the fixed initial state violated) gives the fallback solution.
    The kernel's method of multipliers against the oracle's exact QP solve: same feasible / infeasible verdict,
    first control 1e-4, objective 1e-6 relative, row violation <= 1e-6, slack_used False."""
    from hybrid_controller import workloads as wl
    from hybrid_controller import _native as nat
    B = 96
    x0, xr, ur, ob = wl.config4_instances(B, N=N, seed=40 + N, obstacles=wl.DENSE_OBSTACLES[:M])
    p = nat.mpc_params(kind, N, M, bs, [15, 15, 50], [.1, .1], [30, 30, 40], .3, 5000., 2., 3., .02, soft=False)
    out = emul.mpc(p, x0, xr, ur, ob)
    prm = dict(MPC_PRM20, N=N, block_size=bs, soft=False)
    n_ok = n_fb = 0
    for b in range(B):
        o = (mpc_qp.solve_with_ltv if kind == 0 else mpc_qp.solve)(x0[b], xr[b], ur[b], ob[b], **prm)
        assert (o["status"] == "optimal") == (out["status"][b] == 0), b
        assert np.abs(out["u0"][b] - o["optimal_control"]).max() <= 1e-4 and not out["slack_used"][b]
        if o["status"] == "optimal":
            n_ok += 1
            assert abs(out["cost"][b] - o["cost"]) <= 1e-6 * abs(o["cost"])
            assert np.abs(out["xpred"][b] - o["predicted_states"]).max() <= 1e-5
            # the rows themselves, on the returned prediction: n.(p_k - o) >= d_safe + r to 1e-6
            pref = xr[b][:N, :2] if kind == 0 else np.repeat(xr[b][:1, :2], N, 0) * 0 + xr[b][:N, :2]
            for m in range(M):
                d = pref - ob[b, m, :2]
                dist = np.linalg.norm(d, axis=1)
                keep = dist > 0.01
                lhs = ((out["xpred"][b][:N, :2] - ob[b, m, :2]) * d).sum(1)[keep] / dist[keep]
                assert (lhs >= 0.3 + ob[b, m, 2] - 1e-6).all()
        else:
            n_fb += 1
            assert np.isinf(out["cost"][b]) and out["status"][b] == 3
    assert n_ok >= B // 2 and n_fb >= 1
