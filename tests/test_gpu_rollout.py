"""GPU parity of the closed-loop rollout engine (b2c_rollout_host -> csrc/rollout.cu + K2 + K3) against the shipped
logs, the host emulation of the same per-robot code and the CPU oracle."""
import os

import numpy as np
import pytest

import emul
from conftest import DEFAULT_OBS, MPC_PRM6, assert_g_mpc_closed_loop
from oracle import lin_dare, rollout as oro

pytestmark = pytest.mark.gpu


def _controllers(N=6, block_size=1):
    from hybrid_controller.controllers import LQRController, MPCController, RiskMetrics
    from hybrid_controller import workloads as wl
    return (LQRController(**wl.SIM_LQR), MPCController(horizon=N, block_size=block_size, **wl.SIM_MPC),
            RiskMetrics(d_safe=0.3, d_trigger=1.0, alpha=0.6, beta=0.4, threshold_low=0.2, threshold_medium=0.5))


def test_rollout_lqr_mode_vs_g_lqr(golden):
    from hybrid_controller.rollout import simulate_batch
    g = golden["g_lqr"]
    lqr, _, _ = _controllers()
    table = lin_dare.figure8(20.0)
    out = simulate_batch("lqr", table, [0], table[0:1, 1:4], np.zeros((0, 3)), 998, lqr=lqr, log=True)
    assert np.abs(out["states"][0] - g["states"][:999]).max() < 1e-9           # <= 1e-9 over 999 ticks
    assert np.abs(out["controls"][0] - g["controls"][:998]).max() < 1e-9
    assert out["lqr_steps"][0] == 998


def test_rollout_mpc_mode_vs_g_mpc(golden):
    from hybrid_controller.rollout import simulate_batch
    g = golden["g_mpc"]
    _, mpc, _ = _controllers(N=6, block_size=2)
    table = lin_dare.figure8(20.0)
    out = simulate_batch("mpc", table, [0], table[0:1, 1:4], np.array(DEFAULT_OBS), 998, mpc=mpc, mpc_rate=5, log=True)
    assert_g_mpc_closed_loop(out["states"][0], out["controls"][0], g["states"][:999], g["controls"][:998])
    assert out["mpc_solves"][0] == 200 and out["mpc_fallbacks"][0] == 0
    # collision census of run_mpc_simulation (:288-293): states inside radius + d_safe of any obstacle
    ob = np.array(DEFAULT_OBS)
    inside = (np.linalg.norm(out["states"][0][:, None, :2] - ob[None, :, :2], axis=2) < ob[:, 2] + 0.3).any(axis=1)
    assert out["collision_states"][0] == inside.sum() == (np.linalg.norm(
        g["states"][:999, None, :2] - ob[None, :, :2], axis=2) < ob[:, 2] + 0.3).any(axis=1).sum()


def test_rollout_hybrid_vs_emulation_and_oracle():
    """512 robots x 150 ticks: device == host emulation of the same code (states to 1e-9, identical controller
    choices), and a subset against the independent oracle loop."""
    from hybrid_controller.rollout import simulate_batch
    lqr, mpc, risk = _controllers()
    table = lin_dare.figure8(20.0)
    rng = np.random.default_rng(11)
    R, T = 512, 150
    k0 = rng.integers(0, 900, R).astype(np.int32)
    x0 = table[k0, 1:4] + rng.uniform(-0.08, 0.08, (R, 3))
    obs = np.broadcast_to(np.array(DEFAULT_OBS)[None], (R, 3, 3)).copy()
    obs[:, :, :2] += rng.uniform(-0.2, 0.2, (R, 3, 2))
    sc = (np.arange(R) % 12).astype(np.int32)
    out = simulate_batch("hybrid", table, k0, x0, obs, T, lqr=lqr, mpc=mpc, risk=risk, mpc_step_count=sc, log=True)
    ref = emul.rollout("hybrid", table, k0, x0, obs, T, lqr=lqr, mpc=mpc, risk=risk, mpc_step_count=sc)
    same = (out["controller"] == ref["controller"]).all(axis=1)
    assert same.mean() > 0.99                      # a risk value within rounding of the threshold may flip a robot
    assert np.abs(out["states"][same] - ref["states"][same]).max() < 1e-9
    assert np.abs(out["controls"][same] - ref["controls"][same]).max() < 1e-8
    for key in ("lqr_steps", "mpc_solves", "switches", "collision_states", "mpc_fallbacks", "mpc_step_count"):
        assert np.array_equal(out[key][same], ref[key][same]), key
    assert np.abs(out["mean_error"][same] - ref["mean_error"][same]).max() < 1e-10
    assert 0.05 < out["mpc_solves"].sum() / (R * T) < 0.95 and out["switches"].sum() > R // 4
    prm = dict(MPC_PRM6, block_size=1)
    for r in (0, 17, 300):
        o = oro.simulate("hybrid", table, int(k0[r]), x0[r], obs[r], T, mpc_prm=prm, step_count=int(sc[r]))
        assert np.array_equal(out["controller"][r], o["controller"])
        assert np.abs(out["states"][r] - o["states"]).max() < 1e-6 and np.abs(out["controls"][r] - o["controls"]).max() < 1e-5


def test_rollout_is_sharding_invariant_and_log_free_path():
    """Robot r gives the same result whatever batch it is part of (8e), with and without the per-tick logs."""
    from hybrid_controller.rollout import simulate_batch
    lqr, mpc, risk = _controllers()
    table = lin_dare.figure8(20.0)
    rng = np.random.default_rng(3)
    R, T = 3000, 60
    k0 = rng.integers(0, 950, R).astype(np.int32)
    x0 = table[k0, 1:4] + rng.uniform(-0.08, 0.08, (R, 3))
    obs = np.broadcast_to(np.array(DEFAULT_OBS)[None], (R, 3, 3)).copy()
    obs[:, :, :2] += rng.uniform(-0.2, 0.2, (R, 3, 2))
    full = simulate_batch("hybrid", table, k0, x0, obs, T, lqr=lqr, mpc=mpc, risk=risk)
    cut = 1234
    a = simulate_batch("hybrid", table, k0[:cut], x0[:cut], obs[:cut], T, lqr=lqr, mpc=mpc, risk=risk, log=True)
    b = simulate_batch("hybrid", table, k0[cut:], x0[cut:], obs[cut:], T, lqr=lqr, mpc=mpc, risk=risk)
    for key in ("x", "lqr_steps", "mpc_solves", "switches", "collision_states", "mean_error", "final_error"):
        assert np.array_equal(full[key], np.concatenate([a[key], b[key]])), key
    assert np.array_equal(a["states"][:, -1], a["x"])


def test_rollout_without_obstacles_and_single_robot():
    """No obstacles: the risk is 0, the hybrid loop never leaves LQR and equals the LQR loop; R = 1 works."""
    from hybrid_controller.rollout import simulate_batch
    lqr, mpc, risk = _controllers()
    table = lin_dare.figure8(20.0)
    x0 = table[5:6, 1:4] + 0.03
    a = simulate_batch("hybrid", table, [5], x0, np.zeros((0, 3)), 200, lqr=lqr, mpc=mpc, risk=risk, log=True)
    b = simulate_batch("lqr", table, [5], x0, np.zeros((0, 3)), 200, lqr=lqr, log=True)
    assert a["mpc_solves"][0] == 0 and a["switches"][0] == 0 and a["lqr_steps"][0] == 200
    assert np.array_equal(a["states"], b["states"]) and np.array_equal(a["controls"], b["controls"])
    assert (a["risk"] == 0).all() and a["collision_states"][0] == 0


def test_config5_device_generator_and_shard_invariance():
    """BASELINE config 5 inputs are generated ON THE DEVICE from (seed, global robot index): the device generator equals
    its numpy mirror bit for bit, and a 2-way split of 4,096 robots (two calls, as two ranks would make them) equals the
    unsplit run bit for bit -- states, counters and statistics."""
    from hybrid_controller.rollout import simulate_generated_shard, simulate_batch
    from hybrid_controller import workloads as wl
    lqr, mpc, risk = _controllers()
    table = lin_dare.figure8(20.0)
    R, T, seed = 4096, 60, 5
    kw = dict(lqr=lqr, mpc=mpc, risk=risk)
    full = simulate_generated_shard(R, 0, seed, table, T, return_inputs=True, **kw)
    k0, x0, ob = wl.config5_robots(R, 0, seed, table=table)
    assert np.array_equal(full["inputs"][0], k0) and np.array_equal(full["inputs"][1], x0)
    assert np.array_equal(full["inputs"][2], ob)
    assert k0.min() >= 0 and k0.max() <= 499 and np.abs(x0 - table[k0, 1:4]).max() <= 0.2
    halves = [simulate_generated_shard(R // 2, h * (R // 2), seed, table, T, **kw) for h in range(2)]
    for k in ("x", "lqr_steps", "mpc_solves", "switches", "collision_states", "mpc_fallbacks", "mean_error", "max_error"):
        assert np.array_equal(np.concatenate([h[k] for h in halves]), full[k]), k
    assert full["mpc_solves"].sum() > 0 and full["lqr_steps"].sum() > 0 and full["mpc_fallbacks"].sum() == 0
    # and the engine fed the mirrored inputs through the host entry point gives the same robots
    ref = simulate_batch("hybrid", table, k0[:512], x0[:512], ob[:512], T, **kw)
    assert np.array_equal(ref["x"], full["x"][:512]) and np.array_equal(ref["mpc_solves"], full["mpc_solves"][:512])
