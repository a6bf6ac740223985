#!/usr/bin/env python3
"""Per-warp phase timing of k_mpc_slots (instrumented build: `make -C <pkg>/csrc trace`).

    B2C_LIB=tools/_trace/libb2c_trace.so python tools/slot_trace.py [--B 65536]
Prints, for one launch: when warps finish relative to the first start, rounds per warp, cycles per round in
the backward / refill (finish + setup) / forward phases, and the mean number of busy lanes per round."""
import argparse
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.environ.setdefault("B2C_LIB", os.path.join(ROOT, "tools", "_trace", "libb2c_trace.so"))
sys.path.insert(0, os.path.join(ROOT, "risk-aware-hybrid-lqr-mpc-navigation-for-autonomous-systems_b200"))
import numpy as np  # noqa: E402
import torch  # noqa: E402
from hybrid_controller import _native as nat, workloads as wl  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--B", type=int, default=65536)
    ap.add_argument("--seed", type=int, default=0)
    args = ap.parse_args()
    B, N = args.B, 20
    dev = torch.device("cuda", 0)
    ctx = nat.Context(0)
    reps = -(-B // 65536)
    arrs = [np.concatenate([wl.config4_instances(65536, N=N, seed=args.seed + s)[j] for s in range(reps)])[:B]
            for j in range(4)]
    d = [torch.from_numpy(np.ascontiguousarray(a.reshape(B, -1))).to(dev).t().contiguous() for a in arrs]
    u0 = torch.empty((2, B), dtype=torch.float64, device=dev)
    cost = torch.empty(B, dtype=torch.float64, device=dev)
    st, it, sl = (torch.empty(B, dtype=torch.int32, device=dev) for _ in range(3))
    prm = nat.mpc_params(nat.MPC_LTV, N, 5, 1, [15, 15, 50], [.1, .1], [30, 30, 40], .3, 5000., 2., 3., .02, engine=2)
    for _ in range(4):
        ctx.call("b2c_mpc_dev", B, C.byref(prm), d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), d[3].data_ptr(),
                 None, u0.data_ptr(), cost.data_ptr(), st.data_ptr(), it.data_ptr(), sl.data_ptr(), None, None)
        ctx.sync()
    nw = min(8192, 148 * 8)
    buf = (C.c_ulonglong * (16 * nw))()
    f = ctx.lib.b2c_debug_slot_trace
    f.argtypes = [C.c_void_p, C.c_int]
    assert f(buf, nw) == 0
    t = np.frombuffer(buf, dtype=np.uint64).reshape(nw, 16).astype(np.float64)
    t = t[t[:, 7] == 1]
    t0 = t[:, 0].min()
    end = (t[:, 1] - t0) / 1e3
    rounds = t[:, 2]
    cyc = t[:, 3:6]
    q = lambda a: [float(np.percentile(a, p)) for p in (0, 10, 50, 90, 100)]
    raw = it.cpu().numpy().astype(np.int64)
    it = torch.from_numpy(raw & 0xff)
    res, r0 = (raw >> 8) & 0xff, (raw >> 16) & 0xffff          # rounds resident, warp round of the claim
    itn = it.numpy()
    extra = res - (itn + 1)                                     # rounds beyond one per iteration + the first evaluation
    hard = np.argsort(-(r0 + res))[:12]
    per_instance = {"rounds_resident_pct_50_90_99_100": [float(np.percentile(res, p)) for p in (50, 90, 99, 100)],
                    "extra_rounds_mean": float(extra.mean()), "extra_rounds_max": int(extra.max()),
                    "extra_rounds_hist": np.bincount(np.clip(extra, 0, 12)).tolist(),
                    "last_finishers": [dict(claimed_in_round=int(r0[i]), rounds=int(res[i]), iters=int(itn[i])) for i in hard],
                    "claim_round_of_iters_ge_10": sorted(r0[itn >= 10].tolist())[-10:],
                    "rounds_of_iters_ge_10": sorted(res[itn >= 10].tolist())[-10:]}
    print(json.dumps({
        "per_instance": per_instance,
        "B": B, "warps": len(t), "iters_mean": float(it.float().mean()), "iters_max": int(it.max()),
        "start_spread_us": float((t[:, 0].max() - t0) / 1e3),
        "warp_end_us_pct_0_10_50_90_100": q(end),
        "rounds_per_warp_pct": q(rounds),
        "busy_lanes_per_round": float(t[:, 6].sum() / rounds.sum()),
        "cycles_per_round": {"backward": float(cyc[:, 0].sum() / rounds.sum()),
                             "refill": float(cyc[:, 1].sum() / rounds.sum()),
                             "forward": float(cyc[:, 2].sum() / rounds.sum())},
        "slots_cfg": os.environ.get("B2C_SLOTS_CFG", "auto"),
        "forward_phase_cycles_per_round": dict(zip(("undo", "team_setup", "sweep", "post"), (t[:, 8:12].sum(axis=0) / rounds.sum()).tolist())),
        "team_rounds_frac": float(t[:, 12].sum() / rounds.sum()),
        "forward_cycles_per_team_round": float(t[:, 13].sum() / max(t[:, 12].sum(), 1)),
        "backward_sweeps_per_round": float(t[:, 14].sum() / rounds.sum()),
        "slowest_warps": [dict(end_us=float((r[1] - t0) / 1e3), rounds=int(r[2]), bwd=int(r[3]), refill=int(r[4]), fwd=int(r[5]),
                               undo=int(r[8]), team=int(r[9]), sweep=int(r[10]), post=int(r[11]), team_rounds=int(r[12]),
                               bwd_sweeps=int(r[14])) for r in t[np.argsort(-t[:, 1])[:4]]],
        "us_per_round_mean": float(((t[:, 1] - t[:, 0]) / 1e3).sum() / rounds.sum()),
    }))


if __name__ == "__main__":
    main()
