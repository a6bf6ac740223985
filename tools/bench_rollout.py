#!/usr/bin/env python3
"""Config 5 (BASELINE.json): Monte-Carlo hybrid LQR-MPC closed-loop rollouts, R robots x T ticks on one GPU
(device-resident inputs; robots keyed by their global index so a shard gives the same robots on any GPU count).
Prints one JSON line: robot-steps/s, MPC-solve fraction, switches, collisions, mean tracking error."""
import argparse
import ctypes as C
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "risk-aware-hybrid-lqr-mpc-navigation-for-autonomous-systems_b200"))
import numpy as np  # noqa: E402
import torch  # noqa: E402
from hybrid_controller import _native as nat, workloads as wl, rollout as ro  # noqa: E402
from hybrid_controller.controllers import LQRController, MPCController, RiskMetrics  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--robots", type=int, default=1 << 20)
    ap.add_argument("--ticks", type=int, default=500)
    ap.add_argument("--mode", default="hybrid")
    ap.add_argument("--start", type=int, default=0, help="global index of the first robot of this shard")
    args = ap.parse_args()
    R, T = args.robots, args.ticks
    dev = torch.device("cuda", 0)
    ctx = nat.Context(0)
    table = wl.figure8_table()
    gidx = np.arange(args.start, args.start + R, dtype=np.uint64)
    # counter-based stream keyed by the GLOBAL robot index: SplitMix64 of (index, lane)
    def u01(lane):
        z = (gidx * np.uint64(0x9E3779B97F4A7C15) + np.uint64(lane) * np.uint64(0xBF58476D1CE4E5B9)) & np.uint64(0xFFFFFFFFFFFFFFFF)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
        return (z >> np.uint64(11)).astype(np.float64) / float(1 << 53)
    k0 = np.minimum((u01(0) * 500).astype(np.int32), 499)
    x0 = table[k0, 1:4] + np.stack([(u01(1) - .5) * .2, (u01(2) - .5) * .2, (u01(3) - .5) * .4], 1)
    obs = np.broadcast_to(wl.DEFAULT_OBSTACLES[None], (R, 3, 3)).copy()
    for m in range(3):
        obs[:, m, 0] += (u01(4 + 2 * m) - .5) * .4
        obs[:, m, 1] += (u01(5 + 2 * m) - .5) * .4
    lqr, mpc = LQRController(**wl.SIM_LQR), MPCController(horizon=6, block_size=1, **wl.SIM_MPC)
    risk = RiskMetrics(d_safe=0.3, d_trigger=1.0, alpha=0.6, beta=0.4, threshold_low=0.2, threshold_medium=0.5)
    p = ro.rollout_params(args.mode, T, table.shape[0], lqr, mpc, risk, 3)
    d_tab = torch.from_numpy(table).to(dev)
    d_k0 = torch.from_numpy(k0).to(dev)
    d_obs = torch.from_numpy(np.ascontiguousarray(obs.reshape(R, 9).T)).to(dev)
    x_init = torch.from_numpy(np.ascontiguousarray(x0.T)).to(dev)
    cnt = torch.zeros((5, R), dtype=torch.int32, device=dev)
    st = torch.zeros((3, R), dtype=torch.float64, device=dev)

    def run():
        d_x = x_init.clone()
        torch.cuda.synchronize()
        l0, t0 = ctx.launches, time.perf_counter()
        ctx.call("b2c_rollout_dev", R, C.byref(p), d_tab.data_ptr(), d_k0.data_ptr(), d_x.data_ptr(), d_obs.data_ptr(),
                 None, cnt.data_ptr(), st.data_ptr(), None, None, None, None)
        torch.cuda.synchronize()
        return time.perf_counter() - t0, ctx.launches - l0
    run_small = ro.rollout_params(args.mode, 5, table.shape[0], lqr, mpc, risk, 3)
    ctx.call("b2c_rollout_dev", R, C.byref(run_small), d_tab.data_ptr(), d_k0.data_ptr(), x_init.clone().data_ptr(),
             d_obs.data_ptr(), None, cnt.data_ptr(), st.data_ptr(), None, None, None, None)      # warm-up (allocations)
    dt, launches = run()
    c = cnt.sum(dim=1).cpu().numpy().astype(float)
    print(json.dumps({"metric": "closed-loop robot-steps/s (config 5 shard)", "mode": args.mode, "robots": R, "ticks": T,
                      "seconds": dt, "robot_steps_per_s": R * T / dt, "mpc_solves": int(c[1]),
                      "mpc_solve_fraction": c[1] / (R * T), "mpc_solves_per_s": c[1] / dt, "switches_per_robot": c[2] / R,
                      "collision_state_fraction": c[3] / (R * (T + 1)), "mpc_fallbacks": int(c[4]),
                      "mean_tracking_error_m": float((st[0] / T).mean().item()), "gpu_launches": int(launches)}))


if __name__ == "__main__":
    main()
