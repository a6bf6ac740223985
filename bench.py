#!/usr/bin/env python3
"""Headline benchmark: MPC QP solves/sec (diff-drive, N=20, 5 obstacles) -- BASELINE.json config 4.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
One process per GPU (torchrun for N>1); a "step" = one pass of the K3 kernel over one batch of
65,536 independent solve_with_ltv instances.  Weak scaling: every rank owns its own 65,536-instance
shard of a global batch of N*65,536 (no data-path collective; one all-reduce(MAX) of the time).
Prints ONE JSON line on rank 0.  See DESIGN.md section 6 for how each field is obtained.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "risk-aware-hybrid-lqr-mpc-navigation-for-autonomous-systems_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402

B_PER_GPU = 65536
HORIZON, N_OBS = 20, 5
METRIC = "MPC QP solves/sec (diff-drive, N=20, 5 obstacles)"
WORKLOAD = "config4: 65,536 independent solve_with_ltv instances per GPU, N=20, M=5 dense obstacles, block_size=1"
# algorithmic bytes per solve: inputs 24+504+320+120 = 968 B, outputs u0+cost+status+iters = 32 B
ALG_BYTES_PER_SOLVE = 1000.0
# counted fp64 flops per stage: backward Riccati stage / forward stage with 5 rows
FLOP_BWD_STAGE, FLOP_FWD_STAGE = 262.0, 215.0
ORACLE_PRM = dict(N=HORIZON, Qd=[15, 15, 50], Rd=[.1, .1], Pd=[30, 30, 40], d_safe=.3, rho=5000., v_max=2.,
                  w_max=3., dt=.02, block_size=1)


def _oracle_solve(args):
    from oracle import mpc_qp
    x0, xr, ur, ob = args
    o = mpc_qp.solve_with_ltv(x0, xr, ur, ob, **ORACLE_PRM)
    return o["optimal_control"]


def cpu_arm(n_instances, seed, pool, global_batch=None):
    """Time the CPU implementation of the path on `n_instances` config-4 instances (all host cores).
    With `global_batch` the instances are the FIRST n of that seeded global batch (same as the GPU's)."""
    from hybrid_controller import workloads as wl
    x0, xr, ur, ob = wl.config4_instances(global_batch or n_instances, N=HORIZON, seed=seed, start=0,
                                          count=n_instances)
    items = [(x0[i], xr[i], ur[i], ob[i]) for i in range(n_instances)]
    t0 = time.perf_counter()
    res = pool.map(_oracle_solve, items, chunksize=max(1, n_instances // (4 * pool._processes)))
    dt = time.perf_counter() - t0
    return n_instances / dt, dt, np.array(res)


def _worker_init():
    # one BLAS thread per worker: the pool already uses every core
    try:
        from threadpoolctl import threadpool_limits
        _worker_init.limit = threadpool_limits(1)
    except Exception:
        pass


def make_pool():
    import multiprocessing as mp
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    return mp.get_context("fork").Pool(cores, initializer=_worker_init), cores


def run_reference(args):
    """--impl reference: the reference's CPU path for this metric.  cvxpy/OSQP are absent from the
    image (un-vendored dependency), so the timed code is the oracle port of solve_with_ltv (line-by-
    line QP restatement + HiGHS + KKT polish), on every host core."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    pool, cores = make_pool()
    sample = max(64, 16 * cores)
    for w in range(args.warmup):
        cpu_arm(max(cores, 16), 1000 + w, pool)
    t_total, n_total = 0.0, 0
    for s in range(args.steps):
        _, dt, _ = cpu_arm(sample, 2000 + s, pool)
        t_total += dt
        n_total += sample
    pool.close()
    v = n_total / t_total
    line = {"metric": METRIC, "value": v, "unit": "solves/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * t_total / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "impl": "reference",
            "config": {"workload": WORKLOAD, "sample_per_step": sample},
            "cpu_baseline": {"value": v, "unit": "solves/s", "cores": cores, "kind": "port",
                             "sample": f"{sample} config-4 instances per step x {args.steps} steps; oracle port "
                                       "(HiGHS active-set + KKT polish), cvxpy absent from the image"},
            "e2e": {"value": v, "unit": "solves/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                parts = [p.strip() for p in out.strip().split(",")]
                if len(parts) >= 6:
                    self.rows.append(parts)
            except Exception:
                pass
            time.sleep(0.05)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        sm = sorted(float(r[0]) for r in self.rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for j, n in enumerate(names) if any(r[2 + j].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons,
                "samples": len(self.rows)}


def secondary_metrics(ctx, dev, stream):
    """SURVEY 8(d) secondary figures, device-resident, CUDA-event timed on the context's stream (N=1 only):
    config 3 (2^20 DARE gains, 2^20 linearizations), batch=1 MPC latency at the run_simulation shape (config 2) and a
    config-5 shard of the hybrid closed loop (65,536 robots x 100 ticks)."""
    import ctypes as C
    import torch
    from hybrid_controller import _native as nat, workloads as wl, rollout as ro
    from hybrid_controller.controllers import LQRController, MPCController, RiskMetrics
    out = {}

    def timed(fn, reps):
        fn(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(reps):
            fn()
        e1.record(stream); torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e-3 / reps

    n = 1 << 20
    v, th = wl.config3_points(n)
    dv, dth = torch.from_numpy(v).to(dev), torch.from_numpy(th).to(dev)
    Kd = torch.empty((6, n), dtype=torch.float64, device=dev)
    Pd = torch.empty((6, n), dtype=torch.float64, device=dev)
    it = torch.empty(n, dtype=torch.int32, device=dev); st = torch.empty(n, dtype=torch.int32, device=dev)
    Q, Rr = (C.c_double * 3)(15., 15., 8.), (C.c_double * 2)(.1, .1)
    t = timed(lambda: ctx.call("b2c_dare_dev", n, nat.LIN_EULER, nat.GUARD_LQR, 0.02, Q, Rr, dv.data_ptr(), dth.data_ptr(),
                               Kd.data_ptr(), Pd.data_ptr(), it.data_ptr(), st.data_ptr()), 10)
    flops = 520.0 * float(it.float().mean().item()) + 160.0
    out["dare_config3"] = {"points": n, "gains_per_s": n / t, "ms": t * 1e3, "mean_doubling_steps": float(it.float().mean().item()),
                           "ok_frac": float((st == 0).float().mean().item()), "alg_bytes_per_point": 136,
                           "hbm_gbs": 136 * n / t / 1e9, "fp64_tflops": flops * n / t / 1e12}
    Ad = torch.empty((9, n), dtype=torch.float64, device=dev); Bd = torch.empty((6, n), dtype=torch.float64, device=dev)
    t = timed(lambda: ctx.call("b2c_linearize_dev", n, nat.LIN_EULER, nat.GUARD_NONE, 0.02, dv.data_ptr(), dth.data_ptr(),
                               Ad.data_ptr(), Bd.data_ptr()), 20)
    out["linearize_config3"] = {"points": n, "points_per_s": n / t, "ms": t * 1e3, "alg_bytes_per_point": 136,
                                "hbm_gbs": 136 * n / t / 1e9}
    # batch = 1 latency through the drop-in class (host numpy in/out, wall clock): run_simulation.py --mode mpc shape
    mpc = MPCController(horizon=6, block_size=2, **wl.SIM_MPC)
    table = wl.figure8_table()
    from hybrid_controller.controllers import Obstacle
    obs = [Obstacle(*o) for o in wl.DEFAULT_OBSTACLES.tolist()]
    lat = []
    for k in range(0, 1000, 5):
        xr, ur = table[k:k + 7, 1:4], table[k:k + 7, 4:6]
        if len(xr) < 7:
            break
        t0 = time.perf_counter()
        mpc.solve_with_ltv(xr[0] + 0.01, xr, ur, obs)
        lat.append((time.perf_counter() - t0) * 1e3)
    lat = np.array(lat[5:])
    out["mpc_batch1_latency_ms"] = {"mean": float(lat.mean()), "p50": float(np.percentile(lat, 50)),
                                    "p99": float(np.percentile(lat, 99)), "solves": int(lat.size),
                                    "reference_shipped_log_mean_ms": 82.6}
    # config-5 shard
    R, T = 65536, 100
    rng = np.random.default_rng(0)
    k0 = rng.integers(0, 500, R).astype(np.int32)
    x0 = table[k0, 1:4] + rng.uniform(-1, 1, (R, 3)) * np.array([.1, .1, .2])
    ob = np.broadcast_to(wl.DEFAULT_OBSTACLES[None], (R, 3, 3)).copy()
    ob[:, :, :2] += rng.uniform(-.2, .2, (R, 3, 2))
    p = ro.rollout_params("hybrid", T, table.shape[0], LQRController(**wl.SIM_LQR),
                          MPCController(horizon=6, block_size=1, **wl.SIM_MPC),
                          RiskMetrics(d_safe=0.3, d_trigger=1.0, alpha=0.6, beta=0.4, threshold_low=0.2, threshold_medium=0.5), 3)
    d_tab = torch.from_numpy(table).to(dev); d_k0 = torch.from_numpy(k0).to(dev)
    d_ob = torch.from_numpy(np.ascontiguousarray(ob.reshape(R, 9).T)).to(dev)
    x_init = torch.from_numpy(np.ascontiguousarray(x0.T)).to(dev)
    cnt = torch.zeros((5, R), dtype=torch.int32, device=dev); sts = torch.zeros((3, R), dtype=torch.float64, device=dev)
    best = 1e9
    for _ in range(3):
        d_x = x_init.clone(); torch.cuda.synchronize()
        t0 = time.perf_counter()
        ctx.call("b2c_rollout_dev", R, C.byref(p), d_tab.data_ptr(), d_k0.data_ptr(), d_x.data_ptr(), d_ob.data_ptr(), None,
                 cnt.data_ptr(), sts.data_ptr(), None, None, None, None)
        torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t0)
    c = cnt.sum(dim=1).cpu().numpy().astype(float)
    out["hybrid_rollout_config5_shard"] = {"robots": R, "ticks": T, "robot_steps_per_s": R * T / best,
                                           "mpc_solve_fraction": c[1] / (R * T), "switches_per_robot": c[2] / R,
                                           "collision_state_fraction": c[3] / (R * (T + 1)),
                                           "mean_tracking_error_m": float((sts[0] / T).mean().item()),
                                           "note": "tools/bench_rollout.py runs the full 1,048,576 x 500 shard"}
    return out


def run_gpu(args):
    import ctypes as C
    import torch
    import torch.distributed as dist
    from hybrid_controller import _native as nat, workloads as wl

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B, N, M = B_PER_GPU, HORIZON, N_OBS
    K, W = args.steps, max(args.warmup, 3)

    stream = torch.cuda.Stream(device=dev)          # kernels, copies and timing events all on this stream
    torch.cuda.set_stream(stream)
    ctx = nat.Context(local, stream=stream.cuda_stream)
    prm = nat.mpc_params(nat.MPC_LTV, N, M, 1, [15, 15, 50], [.1, .1], [30, 30, 40], .3, 5000., 2., 3., .02,
                         engine=args.engine)
    kernel_name = "k_mpc<true> (general, HBM scratch)" if args.engine == 1 else "k_mpc_slots (persistent, smem scratch)"

    # ---- inputs: this rank's shard of the global batch, R rotating distinct batches (> L2 in total)
    R = 4
    gB = B * world
    host_sets, dev_sets = [], []
    for r in range(R):
        x0, xr, ur, ob = wl.config4_instances(gB, N=N, seed=100 + r, start=rank * B, count=B)
        host = [torch.from_numpy(np.ascontiguousarray(a)).pin_memory() for a in
                (x0, xr.reshape(B, -1), ur.reshape(B, -1), ob.reshape(B, -1))]
        host_sets.append(host)
        dev_sets.append([h.to(dev).t().contiguous() for h in host])      # SoA [F][B], resident in HBM
    u0 = torch.empty((2, B), dtype=torch.float64, device=dev)
    cost = torch.empty(B, dtype=torch.float64, device=dev)
    st = torch.empty(B, dtype=torch.int32, device=dev)
    it = torch.empty(B, dtype=torch.int32, device=dev)
    sl = torch.empty(B, dtype=torch.int32, device=dev)

    def step_dev(s):
        d = dev_sets[s % R]
        ctx.call("b2c_mpc_dev", B, C.byref(prm), d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(),
                 d[3].data_ptr(), None, u0.data_ptr(), cost.data_ptr(), st.data_ptr(), it.data_ptr(),
                 sl.data_ptr(), None, None)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for s in range(W):
        step_dev(s)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(K + 1)]
    l0 = ctx.launches
    barrier()
    t_wall0 = time.perf_counter()
    ev[0].record(stream)
    for s in range(K):
        step_dev(s)
        ev[s + 1].record(stream)
    barrier()
    t_wall = time.perf_counter() - t_wall0
    launches = ctx.launches - l0
    per_launch_ms = np.array([ev[s].elapsed_time(ev[s + 1]) for s in range(K)])
    total_ms = ev[0].elapsed_time(ev[K])
    iters_mean = float(it.float().mean().item())
    ok_frac = float((st == 0).float().mean().item())
    slack_frac = float(sl.float().mean().item())
    sat_frac = float(((u0[0].abs() >= 2.0 - 1e-9) | (u0[1].abs() >= 3.0 - 1e-9)).float().mean().item())
    if world > 1:
        t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
    value = gB * K / (total_ms * 1e-3)

    # ---- e2e: the C-ABI host entry point, pinned host buffers, H2D + kernel + D2H per step ---------
    h_u0 = torch.empty((B, 2), dtype=torch.float64).pin_memory()
    h_cost = torch.empty(B, dtype=torch.float64).pin_memory()
    h_st, h_it, h_sl = (torch.empty(B, dtype=torch.int32).pin_memory() for _ in range(3))

    def step_host(s):
        h = host_sets[s % R]
        ctx.call("b2c_mpc_host", B, C.byref(prm), h[0].data_ptr(), h[1].data_ptr(), h[2].data_ptr(),
                 h[3].data_ptr(), None, h_u0.data_ptr(), h_cost.data_ptr(), h_st.data_ptr(),
                 h_it.data_ptr(), h_sl.data_ptr(), None, None)

    Ke = max(5, min(K, 50))
    for s in range(3):
        step_host(s)
    barrier()
    t0 = time.perf_counter()
    for s in range(Ke):
        step_host(s)
    barrier()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    sampler.stop_flag = True
    sampler.join(timeout=2)
    e2e_value = gB * Ke / e2e_s
    h2d = B * (3 + 3 * (N + 1) + 2 * N + 3 * M) * 8
    d2h = B * (2 * 8 + 8 + 3 * 4)
    # the e2e result must equal the device-resident result for the same batch
    step_dev((Ke - 1) % R)
    torch.cuda.synchronize()
    assert torch.equal(u0.t().cpu(), h_u0), "e2e and device-resident results differ"

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant (only) kernel -----------------------------------------------------
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak, which = (peaks["hbm_gbs"], "measured") if "hbm_gbs" in peaks else (6650.0, "fallback")
    traffic = None                  # dram bytes per launch of the dominant kernel, from the committed ncu capture
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "r1_slots_traffic.json")))
        if args.engine != 1 and tr.get("instances_per_launch") == B:
            traffic = tr["dram_bytes_per_launch"]
    except Exception:
        pass
    avg_launch_s = float(per_launch_ms.mean()) * 1e-3
    achieved_gbs = ALG_BYTES_PER_SOLVE * B / avg_launch_s / 1e9
    fp64_peak = ctx.fma_peak(64)
    fp32_peak = ctx.fma_peak(32)
    flop_per_solve = N * ((iters_mean + 1.0) * FLOP_BWD_STAGE + (iters_mean + 1.0) * FLOP_FWD_STAGE)
    achieved_tf = flop_per_solve * B / avg_launch_s / 1e12

    # ---- CPU baseline on a bounded sample (rank 0, N=1 only) ---------------------------------------
    cpu = None
    if world == 1 and not args.no_cpu:
        pool, cores = make_pool()
        cpu_arm(max(cores, 16), 7, pool)
        n_s = max(128, 32 * cores)
        v_cpu, dt_cpu, u_cpu = cpu_arm(n_s, 100, pool, global_batch=gB)   # = dev_sets[0][:n_s]
        pool.close()
        step_dev(0)
        torch.cuda.synchronize()
        parity = float(np.abs(u0.t().cpu().numpy()[:n_s] - u_cpu).max())
        cpu = {"value": v_cpu, "unit": "solves/s", "cores": cores, "kind": "port",
               "sample": f"first {n_s} instances of the timed batch ({dt_cpu:.1f} s); oracle port of solve_with_ltv "
                         "(HiGHS + KKT polish); cvxpy/OSQP absent from the image",
               "max_abs_u0_diff_vs_gpu": parity}

    secondary = None
    if world == 1 and not args.no_secondary:
        secondary = secondary_metrics(ctx, dev, stream)

    line = {"metric": METRIC, "value": value, "unit": "solves/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "instances_per_gpu": B, "global_instances": gB,
                       "l2": f"inputs rotate over {R} distinct batches ({R}x{h2d / 1e6:.0f} MB > 126 MB L2); kernel scratch "
                             f"{nat.load_library().b2c_mpc_scratch_bytes(C.byref(prm), B) / 1e6:.0f} MB",
                       "parallelism": f"batch-sharded x{world}, no data-path collective"},
            "p99_batch_latency_ms": float(np.percentile(per_launch_ms, 99)),
            "p50_batch_latency_ms": float(np.percentile(per_launch_ms, 50)),
            "solver": {"mean_newton_iters": iters_mean, "optimal_frac": ok_frac, "slack_active_frac": slack_frac,
                       "saturated_u0_frac": sat_frac},
            "roofline": {"bound": "hbm", "achieved": achieved_gbs, "peak": hbm_peak, "unit": "GB/s",
                         "frac": achieved_gbs / hbm_peak, "traffic": traffic, "peak_source": which,
                         "kernel": kernel_name, "avg_launch_ms": avg_launch_s * 1e3,
                         "alg_bytes_per_solve": ALG_BYTES_PER_SOLVE},
            "fp64": {"achieved_tflops": achieved_tf, "peak_tflops_measured": fp64_peak,
                     "frac": achieved_tf / fp64_peak, "fp32_peak_tflops_measured": fp32_peak,
                     "flop_per_solve_counted": flop_per_solve},
            "e2e": {"value": e2e_value, "unit": "solves/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": Ke, "ms_per_step": 1e3 * e2e_s / Ke, "api": "b2c_mpc_host (C ABI, pinned host buffers)"},
            "gpu_launches": int(launches),
            "clocks": sampler.summary(),
            "wall_s_timed_region": t_wall}
    if cpu is not None:
        line["cpu_baseline"] = cpu
    if secondary is not None:
        line["secondary"] = secondary
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-secondary", action="store_true", help="skip the secondary (DARE / linearize / rollout) figures")
    ap.add_argument("--engine", type=int, default=0, choices=[0, 1, 2],
                    help="K3 kernel: 0 auto (slot kernel for this shape), 1 general kernel, 2 slot kernel")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
