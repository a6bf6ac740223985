#!/usr/bin/env python3
"""Soak of the slot kernel's storage variants on full-size batches: the gains mirror in tensor memory (default), in
global memory (B2C_SLOTS_MIRROR=0) and the 256-slot configuration with the U strips in tensor memory must give
bit-identical outputs on every instance; also reports how many instances had a rejected trial (the only ones
that ever read the mirror).   python tools/soak_mirror.py [--seeds 3] [--B 65536]"""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "risk-aware-hybrid-lqr-mpc-navigation-for-autonomous-systems_b200"))
import numpy as np  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seeds", type=int, default=3)
    ap.add_argument("--B", type=int, default=65536)
    args = ap.parse_args()
    from hybrid_controller.controllers.mpc_controller import MPCController
    from hybrid_controller import workloads as wl
    out = []
    for shape in ((20, 5), (6, 3)):
        N, M = shape
        for seed in range(args.seeds):
            x0, xr, ur, ob = wl.config4_instances(args.B, N=N, seed=100 + seed, obstacles=wl.DENSE_OBSTACLES[:M],
                                                  pos_jitter=0.15 + 0.05 * seed, th_jitter=0.3, obs_jitter=0.25)
            mpc = MPCController(horizon=N, block_size=1, **wl.SIM_MPC)
            mpc.engine = 2
            res = {}
            for name, env in (("tmem_mirror", {}), ("global_mirror", {"B2C_SLOTS_MIRROR": "0"}),
                              ("u_in_tmem_256", {"B2C_SLOTS_CFG": "256,16"})):
                for k in ("B2C_SLOTS_MIRROR", "B2C_SLOTS_CFG"):
                    os.environ.pop(k, None)
                os.environ.update(env)
                res[name] = mpc.solve_with_ltv_batch(x0, xr, ur, ob)
            ref = res["tmem_mirror"]
            same = {n: all(np.array_equal(ref[k], r[k], equal_nan=True) for k in ("u0", "cost", "iters", "status", "useq", "xpred"))
                    for n, r in res.items() if n != "tmem_mirror"}
            out.append(dict(N=N, M=M, seed=seed, B=args.B, identical=same, iters_max=int(ref["iters"].max()),
                            iters_mean=float(ref["iters"].mean()), ok_frac=float((ref["status"] == 0).mean())))
            print(json.dumps(out[-1]), flush=True)
    assert all(all(o["identical"].values()) for o in out), "storage variants disagree"
    print("soak ok")


if __name__ == "__main__":
    main()
