"""Closed-loop rollout engine on the CPU: the oracle restatement of the reference's simulation
loops against the shipped logs, the risk-supervisor mirror against the reference's own outputs, and the per-robot
device code (csrc/rollout.cu, compiled for the host) against the oracle."""
import os

import numpy as np
import pytest

import emul
from conftest import DEFAULT_OBS, GOLDEN, MPC_PRM6, assert_g_mpc_closed_loop
from oracle import lin_dare, rollout as oro


@pytest.fixture(scope="module")
def g_hyb():
    return np.load(os.path.join(GOLDEN, "g_hyb.npz"))


def _controllers(N=6, block_size=1):
    from hybrid_controller.controllers import LQRController, MPCController, RiskMetrics
    from hybrid_controller import workloads as wl
    lqr = LQRController(**wl.SIM_LQR)
    mpc = MPCController(horizon=N, block_size=block_size, **wl.SIM_MPC)
    risk = RiskMetrics(d_safe=0.3, d_trigger=1.0, alpha=0.6, beta=0.4, threshold_low=0.2, threshold_medium=0.5)
    return lqr, mpc, risk


def test_risk_mirror_matches_reference(g_hyb):
    """RiskMetrics mirror == the reference's RiskMetrics on 200 seeded states (bit-exact: same formula, same order)."""
    from hybrid_controller.controllers import RiskMetrics
    rm = RiskMetrics(d_safe=0.3, d_trigger=1.0, alpha=0.6, beta=0.4, threshold_low=0.2, threshold_medium=0.5)
    assert rm.alpha == float(g_hyb["risk_alpha"]) and rm.beta == float(g_hyb["risk_beta"])
    obs = [dict(x=x, y=y, radius=r) for x, y, r in DEFAULT_OBS]
    for i, s in enumerate(g_hyb["risk_states"]):
        a = rm.assess_risk(s, obs)
        ref = g_hyb["risk_out"][i]
        assert [a.distance_risk, a.combined_risk, a.min_obstacle_distance, a.nearest_obstacle_id, float(a.use_mpc)] == ref.tolist()
        assert rm.assess_risk(s, obs, g_hyb["risk_pred"][i]).combined_risk == g_hyb["risk_pred_combined"][i]
        assert oro.distance_risk(s[0], s[1], np.array(DEFAULT_OBS), 0.3, 1.0) == ref[0]
    assert rm.assess_risk(np.zeros(3), []).combined_risk == 0.0


def test_oracle_hybrid_loop_vs_shipped_log(g_hyb):
    """The only hybrid run the reference ships (logs/*_20260208_003249.csv, 565 MPC / 434 LQR ticks) predates the
    final MPC tuning: from the first MPC tick on its controls chatter between +-3 rad/s, which today's
    solve_with_ltv does not produce.  What it does pin is everything up to that point -- risk, hysteresis, the LQR
    law and the plant: ten LQR ticks to 3e-16 and the first switch to MPC at tick 10 -- and the dwell logic (the
    controller tags agree for the first 94 ticks although the MPC controls differ)."""
    table = lin_dare.figure8(20.0)
    prm = dict(MPC_PRM6, block_size=1)
    o = oro.simulate("hybrid", table, 0, table[0, 1:4], np.array(DEFAULT_OBS), 94, mpc_prm=prm)
    assert np.abs(o["states"][:11] - g_hyb["states"][:11]).max() < 1e-12
    assert np.abs(o["controls"][:10] - g_hyb["controls"][:10]).max() < 1e-12
    assert np.array_equal(o["controller"], g_hyb["controller"][:94]) and o["controller"][10] == 1 and o["controller"][9] == 0


def test_emul_rollout_lqr_mode_vs_g_lqr(golden):
    """Config 1 through the rollout engine (per-row gain table from K2, LQR law, plant): 999 ticks vs the shipped log."""
    g = golden["g_lqr"]
    lqr, _, _ = _controllers()
    table = lin_dare.figure8(20.0)
    out = emul.rollout("lqr", table, [0], table[0:1, 1:4], np.zeros((0, 3)), 998, lqr=lqr)
    assert np.abs(out["states"][0] - g["states"][:999]).max() < 1e-9
    assert np.abs(out["controls"][0] - g["controls"][:998]).max() < 1e-9
    assert out["lqr_steps"][0] == 998 and out["mpc_solves"][0] == 0
    assert abs(out["mean_error"][0] - np.linalg.norm(g["errors"][:998, :2], axis=1).mean()) < 1e-9


def test_emul_rollout_mpc_mode_vs_g_mpc(golden):
    """Config 2 through the rollout engine (solve every 5th tick with move blocking 2, hold, ramp): vs the shipped
    OSQP log at the tolerance stated in test_emul_mpc_closed_loop_vs_g_mpc."""
    g = golden["g_mpc"]
    _, mpc, _ = _controllers(N=6, block_size=2)
    table = lin_dare.figure8(20.0)
    out = emul.rollout("mpc", table, [0], table[0:1, 1:4], np.array(DEFAULT_OBS), 998, mpc=mpc, mpc_rate=5)
    assert_g_mpc_closed_loop(out["states"][0], out["controls"][0], g["states"][:999], g["controls"][:998])
    assert out["mpc_solves"][0] == 200 and out["mpc_fallbacks"][0] == 0


def test_emul_rollout_hybrid_vs_oracle():
    """Hybrid loop, robots started at different phases of the figure-8 with jittered obstacles: per-tick states,
    controls, risk and controller choice against oracle/rollout.py (independent QP solver inside)."""
    lqr, mpc, risk = _controllers()
    table = lin_dare.figure8(20.0)
    rng = np.random.default_rng(5)
    R, T = 4, 90
    k0 = np.array([0, 25, 480, 700], np.int32)
    x0 = table[k0, 1:4] + rng.uniform(-0.05, 0.05, (R, 3))
    obs = np.broadcast_to(np.array(DEFAULT_OBS)[None], (R, 3, 3)).copy()
    obs[:, :, :2] += rng.uniform(-0.15, 0.15, (R, 3, 2))
    out = emul.rollout("hybrid", table, k0, x0, obs, T, lqr=lqr, mpc=mpc, risk=risk)
    prm = dict(MPC_PRM6, block_size=1)
    seen = set()
    for r in range(R):
        o = oro.simulate("hybrid", table, int(k0[r]), x0[r], obs[r], T, mpc_prm=prm)
        assert np.array_equal(out["controller"][r], o["controller"])
        assert np.abs(out["risk"][r] - o["risk"]).max() < 1e-12
        assert np.abs(out["states"][r] - o["states"]).max() < 1e-7
        assert np.abs(out["controls"][r] - o["controls"]).max() < 1e-6
        assert out["switches"][r] == o["switches"] and out["collision_states"][r] == o["collision_states"]
        assert out["mpc_step_count"] is None
        assert abs(out["mean_error"][r] - o["errors"].mean()) < 1e-9 and abs(out["final_error"][r] - o["errors"][-1]) < 1e-7
        seen.update(o["controller"].tolist())
    assert seen == {0, 1} and out["switches"].sum() >= 1


def test_emul_rollout_end_of_table_and_counters():
    """Rows are clamped at the end of the table (get_reference_at_index / get_trajectory_segment); per-robot MPC
    ramp counters are carried in and out."""
    lqr, mpc, risk = _controllers()
    table = lin_dare.figure8(2.0)                                  # 100 rows
    obs = np.array([[0.6, 0.55, 0.1]])
    sc = np.array([3, 0], np.int32)
    out = emul.rollout("hybrid", table, [90, 95], table[[90, 95], 1:4], obs, 20, lqr=lqr, mpc=mpc, risk=risk, mpc_step_count=sc)
    o = oro.simulate("hybrid", table, 90, table[90, 1:4], obs, 20, mpc_prm=dict(MPC_PRM6, block_size=1), step_count=3)
    assert np.abs(out["states"][0] - o["states"]).max() < 1e-7 and np.array_equal(out["controller"][0], o["controller"])
    assert out["mpc_step_count"][0] == o["step_count"] and out["mpc_step_count"][1] == out["mpc_solves"][1]


def test_export_robot_csv_matches_reference_log_schema(golden, tmp_path):
    """One robot of a batch written in the SimulationLogger CSV schema: header identical to the shipped logs, rows
    equal to G-LQR's (reference columns exactly, states / errors to the closed-loop tolerance)."""
    import csv
    from hybrid_controller.rollout import export_robot_csv
    g = golden["g_lqr"]
    lqr, _, _ = _controllers()
    table = lin_dare.figure8(20.0)
    out = emul.rollout("lqr", table, [0, 3], table[[0, 3], 1:4], np.zeros((0, 3)), 998, lqr=lqr)
    ps, pc = export_robot_csv(out, 0, table, 0, str(tmp_path), "test", mode="lqr")
    rows = list(csv.reader(open(ps)))
    assert rows[0] == ["timestep", "px", "py", "theta", "px_ref", "py_ref", "theta_ref", "error_px", "error_py",
                       "error_theta", "error_norm"]
    a = np.array([[float(v) for v in r] for r in rows[1:]])
    assert len(a) == 998 and np.array_equal(a[:, 0], np.arange(998))
    assert np.abs(a[:, 1:4] - g["states"][:998]).max() < 1e-9 and np.abs(a[:, 4:7] - g["refs"][:998]).max() < 1e-12
    assert np.abs(a[:, 7:10] - g["errors"][:998]).max() < 1e-9
    assert np.abs(a[:, 10] - np.linalg.norm(g["errors"][:998], axis=1)).max() < 1e-9
    crow = list(csv.reader(open(pc)))
    assert crow[0] == ["timestep", "v", "omega", "controller", "solve_time_ms"] and crow[1][3] == "LQR" and crow[1][4] == ""
    assert np.abs(np.array([[float(r[1]), float(r[2])] for r in crow[1:]]) - g["controls"][:998]).max() < 1e-9
