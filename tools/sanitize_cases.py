#!/usr/bin/env python3
"""Small invocations of every kernel of libb2c.so for compute-sanitizer (tools/sanitize.sh runs each case under
memcheck / racecheck / synccheck / initcheck).  numpy + ctypes only (no torch: the sanitizer then instruments nothing
but libb2c's own kernels).  Each case also checks its result, so a sanitizer-clean but wrong run does not pass."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "risk-aware-hybrid-lqr-mpc-navigation-for-autonomous-systems_b200"),
          os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np  # noqa: E402


def case_lin_dare():
    from hybrid_controller.models.linearization import Linearizer
    from hybrid_controller.controllers.lqr_controller import LQRController
    from hybrid_controller import workloads as wl
    rng = np.random.default_rng(0)
    v, th = rng.uniform(-2, 2, 257), rng.uniform(-3, 3, 257)
    lin = Linearizer(dt=0.02)
    for meth in ("euler", "exact", "jacobian"):
        lin.get_discrete_model_batch(v, th, meth)
    lin.get_discrete_model_explicit_batch(v, th)
    lqr = LQRController(**wl.SIM_LQR)
    K, P, st = lqr.compute_gain_batch(np.abs(v) + 0.01, th, want_P=True)
    assert (st == 0).all()
    lqr.compute_gain(1.0, 0.3)


def _mpc(N, M, B, engine, bs=1, soft=True, deep=True, seq=True):
    from hybrid_controller.controllers.mpc_controller import MPCController
    from hybrid_controller import workloads as wl
    kw = dict(pos_jitter=0.35, th_jitter=0.6, obs_jitter=0.3) if deep else {}
    x0, xr, ur, ob = wl.config4_instances(B, N=N, seed=5 * N + M, obstacles=wl.DENSE_OBSTACLES[:M], **kw)
    mpc = MPCController(horizon=N, block_size=bs, **wl.SIM_MPC)
    mpc.engine = engine
    sc = (np.arange(B) % 13).astype(np.int32)
    out = mpc.solve_with_ltv_batch(x0, xr, ur, ob, step_count=sc, use_soft_constraints=soft, want_seq=seq)
    assert (out["status"] == 0).mean() > 0.5
    return out


def case_slots_tail():      # helper lanes active from the first round (test_mpc_slot_kernel_helper_lanes... shapes)
    _mpc(20, 5, 3000, 2)
    _mpc(6, 3, 700, 2)


def case_slots_refill():    # refill + prepare + claim order
    _mpc(20, 5, 40000, 2, deep=False, seq=False)


def case_general():         # k_mpc<0/1>: move blocking, LTI, hard rows
    _mpc(20, 5, 600, 1, bs=3)
    _mpc(6, 3, 300, 1, bs=2)
    _mpc(20, 5, 256, 1, soft=False)
    from hybrid_controller.controllers.mpc_controller import MPCController, Obstacle
    from hybrid_controller import workloads as wl
    mpc = MPCController(horizon=10, **wl.SIM_MPC)
    t = wl.figure8_table()
    mpc.solve(t[40, 1:4] + 0.02, t[40:51, 1:4], t[40:50, 4:6], [Obstacle(*o) for o in wl.DEFAULT_OBSTACLES.tolist()])


def case_rollout():
    from hybrid_controller.rollout import simulate_batch
    from hybrid_controller.controllers import LQRController, MPCController, RiskMetrics
    from hybrid_controller import workloads as wl
    table = wl.figure8_table()
    rng = np.random.default_rng(11)
    R, T = 512, 50
    k0 = rng.integers(0, 900, R).astype(np.int32)
    x0 = table[k0, 1:4] + rng.uniform(-0.08, 0.08, (R, 3))
    obs = np.broadcast_to(wl.DEFAULT_OBSTACLES[None], (R, 3, 3)).copy()
    obs[:, :, :2] += rng.uniform(-0.2, 0.2, (R, 3, 2))
    out = simulate_batch("hybrid", table, k0, x0, obs, T, lqr=LQRController(**wl.SIM_LQR),
                         mpc=MPCController(horizon=6, block_size=1, **wl.SIM_MPC),
                         risk=RiskMetrics(d_safe=0.3, d_trigger=1.0, alpha=0.6, beta=0.4, threshold_low=0.2,
                                          threshold_medium=0.5), log=True)
    assert out["mpc_solves"].sum() > 0 and out["lqr_steps"].sum() > 0


CASES = {k[5:]: v for k, v in list(globals().items()) if k.startswith("case_")}

if __name__ == "__main__":
    names = sys.argv[1:] or list(CASES)
    for n in names:
        CASES[n]()
        print("case ok:", n, flush=True)
