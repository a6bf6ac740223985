#!/usr/bin/env python3
"""Batch-size sweep of the K3 kernels (B in {1, 256, 4096, 65536, 1048576}), device-resident
inputs, CUDA-event timing on the launching stream.  Prints one JSON line per (engine, B)."""
import argparse
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "risk-aware-hybrid-lqr-mpc-navigation-for-autonomous-systems_b200"))
import numpy as np  # noqa: E402
import torch  # noqa: E402
from hybrid_controller import _native as nat, workloads as wl  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", default="1,256,4096,65536,1048576")
    ap.add_argument("--engines", default="2,1")
    ap.add_argument("--reps", type=int, default=20)
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    ctx = nat.Context(0, stream=stream.cuda_stream)
    N, M = 20, 5
    sizes = [int(x) for x in args.sizes.split(",")]
    base = wl.config4_instances(65536, N=N, seed=0)
    for B in sizes:
        reps = -(-B // 65536)
        arrs = [np.concatenate([a] * reps)[:B] if B > 65536 else a[:B] for a in base]
        if B > 65536:           # distinct instances: re-draw with another seed per 65,536 block
            arrs = [np.concatenate([wl.config4_instances(65536, N=N, seed=s)[j] for s in range(reps)])[:B]
                    for j in range(4)]
        d = [torch.from_numpy(np.ascontiguousarray(a.reshape(B, -1))).to(dev).t().contiguous() for a in arrs]
        u0 = torch.empty((2, B), dtype=torch.float64, device=dev)
        cost = torch.empty(B, dtype=torch.float64, device=dev)
        st, it, sl = (torch.empty(B, dtype=torch.int32, device=dev) for _ in range(3))
        for eng in [int(e) for e in args.engines.split(",")]:
            prm = nat.mpc_params(nat.MPC_LTV, N, M, 1, [15, 15, 50], [.1, .1], [30, 30, 40], .3, 5000., 2., 3.,
                                 .02, engine=eng)

            def go():
                ctx.call("b2c_mpc_dev", B, C.byref(prm), d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(),
                         d[3].data_ptr(), None, u0.data_ptr(), cost.data_ptr(), st.data_ptr(), it.data_ptr(),
                         sl.data_ptr(), None, None)
            for _ in range(3):
                go()
            torch.cuda.synchronize()
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.reps + 1)]
            ev[0].record(stream)
            for r in range(args.reps):
                go()
                ev[r + 1].record(stream)
            torch.cuda.synchronize()
            ms = np.array([ev[r].elapsed_time(ev[r + 1]) for r in range(args.reps)])
            print(json.dumps({"engine": eng, "B": B, "ms_median": float(np.median(ms)), "ms_p99": float(np.percentile(ms, 99)),
                              "solves_per_s": B / (float(np.median(ms)) * 1e-3), "iters_mean": float(it.float().mean()),
                              "iters_max": int(it.max()), "ok": float((st == 0).float().mean())}), flush=True)


if __name__ == "__main__":
    main()
