#!/usr/bin/env python3
"""Headline benchmark: MPC QP solves/sec (diff-drive, N=20, 5 obstacles) -- BASELINE.json config 4.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
One process per GPU (torchrun for N>1); a "step" = one pass of the K3 kernels over one batch of
65,536 independent solve_with_ltv instances.  Weak scaling: every rank owns its own 65,536-instance
shard of a global batch of N*65,536 (no data-path collective; one all-reduce(MAX) of the time, and the
<= 1 KB statistics vector of the config-5 leg over NCCL).
Prints ONE JSON line on rank 0.  See DESIGN.md section 6 for how each field is obtained.
"""
import argparse
import hashlib
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
# algorithmic bytes per solve: inputs 24+504+320+120 = 968 B, outputs u0+cost+status+iters = 32 B;
# with control_sequence + predicted_states written: + 320 + 504 = 1,824 B
ALG_BYTES_PER_SOLVE, ALG_BYTES_FULL = 1000.0, 1824.0
# counted fp64 flops per stage: backward Riccati stage / forward stage with 5 rows
FLOP_BWD_STAGE, FLOP_FWD_STAGE = 262.0, 215.0
ORACLE_PRM = dict(N=HORIZON, Qd=[15, 15, 50], Rd=[.1, .1], Pd=[30, 30, 40], d_safe=.3, rho=5000., v_max=2.,
                  w_max=3., dt=.02, block_size=1)
KERNEL_SRCS = ("mpc_slots.cu", "mpc_args.cuh", "b2c_common.cuh")


def kernel_src_sha():
    """hash of the sources of the dominant kernel: ties a committed ncu traffic figure to the code it was taken from"""
    h = hashlib.sha256()
    for f in KERNEL_SRCS:
        h.update(open(os.path.join(PKG, "csrc", f), "rb").read())
    return h.hexdigest()[:16]


def _oracle_solve(args):
    from oracle import mpc_qp
    x0, xr, ur, ob = args
    o = mpc_qp.solve_with_ltv(x0, xr, ur, ob, **ORACLE_PRM)
    return o["optimal_control"]


def _oracle_gain(args):
    from oracle import lin_dare
    v, th = args
    return lin_dare.lqr_gain(v, th, [15, 15, 8], [.1, .1], .02)[0]


def cpu_arm(n_instances, seed, pool, global_batch=None):
    """Time the CPU implementation of the path on `n_instances` config-4 instances (all host cores, or this
    process alone when pool is None).  With `global_batch` the instances are the FIRST n of that seeded global
    batch (same as the GPU's)."""
    from hybrid_controller import workloads as wl
    x0, xr, ur, ob = wl.config4_instances(global_batch or n_instances, N=HORIZON, seed=seed, start=0,
                                          count=n_instances)
    items = [(x0[i], xr[i], ur[i], ob[i]) for i in range(n_instances)]
    t0 = time.perf_counter()
    if pool is None:
        res = [_oracle_solve(it) for it in items]
    else:
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
                                       "(HiGHS active-set + KKT polish), cvxpy absent from the image",
                             "reference_shipped_log_ms_per_solve": 82.6},
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


def secondary_metrics(ctx, dev, stream, with_cpu):
    """SURVEY 8(d) secondary figures (N=1 only), device-resident, CUDA-event timed on the context's stream: config 3
    (2^20 DARE gains, 2^20 linearizations, each with its CPU baseline) and batch=1 MPC latency at the
    run_simulation shape (config 2)."""
    import ctypes as C
    import torch
    from hybrid_controller import _native as nat, workloads as wl
    from hybrid_controller.controllers import MPCController
    out = {}

    def timed(fn, reps):
        fn(0); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for r in range(reps):
            fn(r)
        e1.record(stream); torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e-3 / reps

    n, R = 1 << 20, 8                                # R rotating input / output sets: 8 x 143 MB > the 126 MB L2
    v, th = wl.config3_points(n)
    dv = [torch.from_numpy(np.roll(v, 1000 * r)).to(dev) for r in range(R)]
    dth = [torch.from_numpy(np.roll(th, 1000 * r)).to(dev) for r in range(R)]
    Kd = [torch.empty((6, n), dtype=torch.float64, device=dev) for _ in range(R)]
    Pd = [torch.empty((6, n), dtype=torch.float64, device=dev) for _ in range(R)]
    it = torch.empty(n, dtype=torch.int32, device=dev); st = torch.empty(n, dtype=torch.int32, device=dev)
    Q, Rr = (C.c_double * 3)(15., 15., 8.), (C.c_double * 2)(.1, .1)
    t = timed(lambda r: ctx.call("b2c_dare_dev", n, nat.LIN_EULER, nat.GUARD_LQR, 0.02, Q, Rr, dv[r % R].data_ptr(),
                                 dth[r % R].data_ptr(), Kd[r % R].data_ptr(), Pd[r % R].data_ptr(), it.data_ptr(),
                                 st.data_ptr()), 16)
    I = float(it.float().mean().item())
    out["dare_config3"] = {"points": n, "gains_per_s": n / t, "ms": t * 1e3, "mean_doubling_steps": I,
                           "ok_frac": float((st == 0).float().mean().item()), "alg_bytes_per_point": 136,
                           "hbm_gbs": 136 * n / t / 1e9,
                           "fp64_tflops": (470.0 * I + 160.0) * n / t / 1e12,
                           "fp64_tflops_baseline_md_formula": (420.0 * I + 160.0) * n / t / 1e12,
                           "flop_formula": "470*I+160 counted from csrc/lin_dare.cu (five general 3x3 products and three symmetric-"
                                           "result ones per doubling step); BASELINE.md section 4 estimates 420*I+160 -- both reported"}
    Ad = [torch.empty((9, n), dtype=torch.float64, device=dev) for _ in range(R)]
    Bd = [torch.empty((6, n), dtype=torch.float64, device=dev) for _ in range(R)]
    t = timed(lambda r: ctx.call("b2c_linearize_dev", n, nat.LIN_EULER, nat.GUARD_NONE, 0.02, dv[r % R].data_ptr(),
                                 dth[r % R].data_ptr(), Ad[r % R].data_ptr(), Bd[r % R].data_ptr()), 32)
    out["linearize_config3"] = {"points": n, "points_per_s": n / t, "ms": t * 1e3, "alg_bytes_per_point": 136,
                                "hbm_gbs": 136 * n / t / 1e9,
                                "l2": f"inputs and outputs rotate over {R} distinct sets ({R}x{136 * n / 1e6:.0f} MB > 126 MB L2)"}
    del Ad, Bd, Kd, Pd
    if with_cpu:
        # CPU baselines: the reference's own arithmetic -- scipy.linalg.solve_discrete_are + the gain
        # formula of lqr_controller.py:126-132 (oracle/lin_dare.lqr_gain), 4,096-point slice, 1 core and all cores;
        # get_discrete_model_explicit (linearization.py:190-225) as a Python loop over 65,536 points, 1 core
        from oracle import lin_dare
        pts = list(zip(v[:4096].tolist(), th[:4096].tolist()))
        t0 = time.perf_counter()
        for a in pts[:1024]:
            _oracle_gain(a)
        one = 1024 / (time.perf_counter() - t0)
        pool, cores = make_pool()
        pool.map(_oracle_gain, pts[:cores * 4])
        t0 = time.perf_counter()
        pool.map(_oracle_gain, pts, chunksize=64)
        allc = 4096 / (time.perf_counter() - t0)
        pool.close()
        out["dare_config3"]["cpu_baseline"] = {"gains_per_s_1_core": one, "gains_per_s_all_cores": allc, "cores": cores,
                                               "kind": "reference arithmetic (scipy.linalg.solve_discrete_are)",
                                               "sample": "1,024 points on 1 core; 4,096 points on all cores"}
        t0 = time.perf_counter()
        for a, b in zip(v[:65536].tolist(), th[:65536].tolist()):
            lin_dare.discrete_explicit(a, b, 0.02)
        out["linearize_config3"]["cpu_baseline"] = {"points_per_s_1_core": 65536 / (time.perf_counter() - t0), "cores": 1,
                                                    "kind": "port of get_discrete_model_explicit, Python loop",
                                                    "sample": "65,536 points"}
    # the throughput regime of K3: 1,048,576 instances per launch (16 x the headline batch): the slot kernel then runs
    # 256 slots per SM with the U strips in tensor memory (two warps per scheduler) and latency no longer decides
    nb, blocks = 1 << 20, 16
    prm = nat.mpc_params(nat.MPC_LTV, HORIZON, N_OBS, 1, [15, 15, 50], [.1, .1], [30, 30, 40], .3, 5000., 2., 3., .02)
    parts = [wl.config4_instances(B_PER_GPU, N=HORIZON, seed=500 + b) for b in range(blocks)]
    big = [torch.from_numpy(np.ascontiguousarray(np.concatenate([p_[j] for p_ in parts]).reshape(nb, -1))).to(dev).t().contiguous()
           for j in range(4)]
    del parts
    bu0 = torch.empty((2, nb), dtype=torch.float64, device=dev); bcost = torch.empty(nb, dtype=torch.float64, device=dev)
    bst, bit, bsl = (torch.empty(nb, dtype=torch.int32, device=dev) for _ in range(3))
    t = timed(lambda r: ctx.call("b2c_mpc_dev", nb, C.byref(prm), big[0].data_ptr(), big[1].data_ptr(), big[2].data_ptr(),
                                 big[3].data_ptr(), None, bu0.data_ptr(), bcost.data_ptr(), bst.data_ptr(), bit.data_ptr(),
                                 bsl.data_ptr(), None, None), 10)
    I1 = float(bit.float().mean().item())
    flop1 = HORIZON * (I1 + 1.0) * (FLOP_BWD_STAGE + FLOP_FWD_STAGE)
    peak64 = ctx.fma_peak(64)
    out["mpc_1M_instances_per_launch"] = {"instances": nb, "ms": t * 1e3, "solves_per_s": nb / t, "mean_newton_iters": I1,
                                          "optimal_frac": float((bst == 0).float().mean().item()),
                                          "fp64_tflops": flop1 * nb / t / 1e12, "fp64_frac": flop1 * nb / t / 1e12 / peak64,
                                          "hbm_gbs_algorithmic": ALG_BYTES_PER_SOLVE * nb / t / 1e9,
                                          "kernel": "k_mpc_slots<256, 5, 16>: 256 slots per SM, first 16 stages of the U strips in tensor memory",
                                          "l2": "1.0 GB of inputs per launch > 126 MB L2"}
    del big, bu0, bcost, bst, bit, bsl
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
    return out


def config5_leg(ctx, dev, rank, world, robots, ticks):
    """BASELINE config 5: Monte-Carlo hybrid LQR-MPC closed-loop rollouts, `robots` per GPU x `ticks` plant steps,
    batch-sharded by global robot index (inputs generated on the device from (seed, global index)); the aggregates are
    combined with the design's only collective, the <= 1 KB statistics all-reduce (NCCL when world > 1)."""
    import torch
    import torch.distributed as dist
    from hybrid_controller import workloads as wl, dist_stats
    from hybrid_controller.rollout import simulate_generated_shard
    from hybrid_controller.controllers import LQRController, MPCController, RiskMetrics
    kw = dict(lqr=LQRController(**wl.SIM_LQR), mpc=MPCController(horizon=6, block_size=1, **wl.SIM_MPC),
              risk=RiskMetrics(d_safe=0.3, d_trigger=1.0, alpha=0.6, beta=0.4, threshold_low=0.2, threshold_medium=0.5),
              ctx=ctx, device=dev.index)
    table = wl.figure8_table()
    # warm-up at the full robot count (the context's arenas reach their final size, modules are loaded), few ticks
    simulate_generated_shard(robots, rank * robots, 1, table, min(ticks, 10), **kw)
    reps = []
    for _ in range(2):            # two timed repetitions of the same shard (deterministic: identical results)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        out = simulate_generated_shard(robots, rank * robots, 0, table, ticks, **kw)
        el = torch.tensor([out["elapsed_s"]], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(el, op=dist.ReduceOp.MAX)
        reps.append(float(el.item()))
    st = dist_stats.reduce_stats(dist_stats.rollout_stats(out, ticks), device=dev)
    elapsed = float(np.mean(reps))
    steps = float(st["robot_steps"])
    return {"workload": f"config5: {robots:,} robots per GPU x {ticks} ticks, hybrid loop (risk -> hysteresis -> LQR | MPC(N=6, "
                        "block 1) -> Euler plant), default 3 obstacles jittered, inputs generated on the device",
            "robots_global": int(st["robots"]), "ticks": ticks, "elapsed_s_max_over_ranks": elapsed,
            "elapsed_s_each_repetition": reps,
            "robot_steps_per_s": steps / elapsed, "mpc_solves_per_s": float(st["mpc_solves"]) / elapsed,
            "mpc_solve_fraction": float(st["mpc_solves"]) / steps, "switches_per_robot": float(st["switches"]) / float(st["robots"]),
            "collision_state_fraction": float(st["collision_states"]) / (float(st["robots"]) * (ticks + 1)),
            "mpc_fallbacks": int(st["mpc_fallbacks"]),
            "mean_tracking_error_m": float(st["err_mean_sum"]) / float(st["robots"]), "max_tracking_error_m": float(st["err_max"]),
            "stats_reduce": "torch.distributed all_reduce(SUM) + all_reduce(MAX) over " + ("NCCL" if world > 1 else "one rank (no-op)"),
            "scaling": "weak"}


def bind_to_gpu_numa_node(index):
    """Pin this rank to the CPUs next to its GPU (NVML's affinity mask) BEFORE it allocates its pinned host buffers:
    with eight ranks on one box the end-to-end figure is bound by host-side copy bandwidth, and first-touch pinned
    memory on the wrong socket halves it for half the GPUs.  Returns a description for the JSON line."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = {64 * w + b for w, word in enumerate(words) for b in range(64) if (word >> b) & 1}
        allowed = cpus & set(os.sched_getaffinity(0))
        if not allowed:
            return "NVML affinity mask empty inside this cpuset: not bound"
        os.sched_setaffinity(0, allowed)
        return f"bound to {len(allowed)} CPUs of GPU {index}'s NUMA node ({min(allowed)}-{max(allowed)})"
    except Exception as e:                                   # measurement convenience only
        return f"not bound ({type(e).__name__})"


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
    numa = bind_to_gpu_numa_node(local) if world > 1 else "not bound (single rank)"
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B, N, M = B_PER_GPU, HORIZON, N_OBS
    K, W = args.steps, max(args.warmup, 3)

    stream = torch.cuda.Stream(device=dev)          # kernels, copies and timing events all on this stream
    torch.cuda.set_stream(stream)
    ctx = nat.Context(local, stream=stream.cuda_stream)
    prm = nat.mpc_params(nat.MPC_LTV, N, M, 1, [15, 15, 50], [.1, .1], [30, 30, 40], .3, 5000., 2., 3., .02,
                         engine=args.engine)
    kernel_name = "k_mpc<true> (general, HBM scratch)" if args.engine == 1 else \
        "k_mpc_slots (persistent; strips in shared memory" + (", U in tensor memory)" if B >= 4 * 148 * 256 else ")")

    # ---- inputs: this rank's shard of the global batch, R rotating distinct batches (> L2 in total)
    R = 4
    gB = B * world
    table = wl.figure8_table()
    h_table = torch.from_numpy(np.ascontiguousarray(table)).pin_memory()
    host_sets, dev_sets, tab_sets = [], [], []
    for r in range(R):
        x0, xr, ur, ob, k0 = wl.config4_instances(gB, N=N, seed=100 + r, start=rank * B, count=B, return_k0=True)
        host = [torch.from_numpy(np.ascontiguousarray(a)).pin_memory() for a in
                (x0, xr.reshape(B, -1), ur.reshape(B, -1), ob.reshape(B, -1))]
        host_sets.append(host)
        tab_sets.append(torch.from_numpy(np.ascontiguousarray(k0.astype(np.int32))).pin_memory())
        dev_sets.append([h.to(dev).t().contiguous() for h in host])      # SoA [F][B], resident in HBM
    u0 = torch.empty((2, B), dtype=torch.float64, device=dev)
    cost = torch.empty(B, dtype=torch.float64, device=dev)
    st = torch.empty(B, dtype=torch.int32, device=dev)
    it = torch.empty(B, dtype=torch.int32, device=dev)
    sl = torch.empty(B, dtype=torch.int32, device=dev)
    useq = torch.empty((2 * N, B), dtype=torch.float64, device=dev)
    xpred = torch.empty((3 * (N + 1), B), dtype=torch.float64, device=dev)
    scnt = torch.full((B,), 10, dtype=torch.int32, device=dev)

    def step_dev(s, full=False):
        d = dev_sets[s % R]
        ctx.call("b2c_mpc_dev", B, C.byref(prm), d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(),
                 d[3].data_ptr(), scnt.data_ptr() if full else None, u0.data_ptr(), cost.data_ptr(), st.data_ptr(),
                 it.data_ptr(), sl.data_ptr(), useq.data_ptr() if full else None, xpred.data_ptr() if full else None)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(x):
        if world > 1:
            t = torch.tensor([x], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        return x

    def timed_dev(full, steps):
        for s in range(W):
            step_dev(s, full)
        barrier()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
        l0 = ctx.launches
        barrier()
        t0 = time.perf_counter()
        ev[0].record(stream)
        for s in range(steps):
            step_dev(s, full)
            ev[s + 1].record(stream)
        barrier()
        wall = time.perf_counter() - t0
        per = np.array([ev[s].elapsed_time(ev[s + 1]) for s in range(steps)])
        return allmax(ev[0].elapsed_time(ev[steps])), per, ctx.launches - l0, wall

    sampler = ClockSampler(local)
    sampler.start()
    total_ms, per_launch_ms, launches, t_wall = timed_dev(False, K)
    iters_mean = float(it.float().mean().item())
    ok_frac = float((st == 0).float().mean().item())
    slack_frac = float(sl.float().mean().item())
    sat_frac = float(((u0[0].abs() >= 2.0 - 1e-9) | (u0[1].abs() >= 3.0 - 1e-9)).float().mean().item())
    value = gB * K / (total_ms * 1e-3)
    Kf = max(5, min(K, 50))
    full_ms, _, _, _ = timed_dev(True, Kf)
    value_full = gB * Kf / (full_ms * 1e-3)

    # ---- e2e: the C-ABI host entry points, pinned host buffers, H2D + kernels + D2H per step -------------------
    # (a) table form, asynchronous (two calls in flight): each instance's reference segment is
    #     get_trajectory_segment(k0, N+1) of ONE shared table -- what every caller of the reference does before
    #     solve_with_ltv (run_simulation.py:247) -- gathered on the device: 148 B per instance cross PCIe;
    # (b) the same, synchronous; (c) materialised segments (968 B per instance), synchronous, as in round 1.
    outs = [dict(u0=torch.empty((B, 2), dtype=torch.float64).pin_memory(), cost=torch.empty(B, dtype=torch.float64).pin_memory(),
                 st=torch.empty(B, dtype=torch.int32).pin_memory(), it=torch.empty(B, dtype=torch.int32).pin_memory(),
                 sl=torch.empty(B, dtype=torch.int32).pin_memory()) for _ in range(2)]

    def step_table(s, name):
        h, o = host_sets[s % R], outs[s % 2]
        ctx.call(name, B, C.byref(prm), h_table.data_ptr(), table.shape[0], tab_sets[s % R].data_ptr(), h[0].data_ptr(),
                 h[3].data_ptr(), None, o["u0"].data_ptr(), o["cost"].data_ptr(), o["st"].data_ptr(), o["it"].data_ptr(),
                 o["sl"].data_ptr(), None, None)

    def step_host(s, name="b2c_mpc_host"):
        h, o = host_sets[s % R], outs[s % 2]
        ctx.call(name, B, C.byref(prm), h[0].data_ptr(), h[1].data_ptr(), h[2].data_ptr(),
                 h[3].data_ptr(), None, o["u0"].data_ptr(), o["cost"].data_ptr(), o["st"].data_ptr(),
                 o["it"].data_ptr(), o["sl"].data_ptr(), None, None)

    def timed_host(fn, steps, wait):
        for s in range(3):
            fn(s)
        if wait:
            ctx.call("b2c_wait")
        barrier()
        t0 = time.perf_counter()
        for s in range(steps):
            fn(s)
        if wait:
            ctx.call("b2c_wait")
        barrier()
        return allmax(time.perf_counter() - t0)

    Ke = max(5, min(K, 50))
    e2e_async_s = timed_host(lambda s: step_table(s, "b2c_mpc_table_host_async"), Ke, True)
    e2e_sync_s = timed_host(lambda s: step_table(s, "b2c_mpc_table_host"), Ke, False)
    last = (Ke - 1)
    res_table = outs[last % 2]["u0"].clone()
    e2e_mat_async_s = timed_host(lambda s: step_host(s, "b2c_mpc_host_async"), Ke, True)
    e2e_mat_s = timed_host(step_host, Ke, False)
    sampler.stop_flag = True
    sampler.join(timeout=2)
    h2d_table = B * (3 * 8 + 3 * M * 8 + 4) + table.size * 8
    h2d_mat = B * (3 + 3 * (N + 1) + 2 * N + 3 * M) * 8
    d2h = B * (2 * 8 + 8 + 3 * 4)
    # every path must return the device-resident result for the same batch, bit for bit
    step_dev(last % R)
    torch.cuda.synchronize()
    assert torch.equal(u0.t().cpu(), outs[last % 2]["u0"]), "materialised e2e and device-resident results differ"
    assert torch.equal(u0.t().cpu(), res_table), "table-form e2e and device-resident results differ"

    c5 = None
    if not args.no_config5:
        c5 = config5_leg(ctx, dev, rank, world, args.config5_robots, args.config5_ticks)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- rooflines of the dominant kernel ------------------------------------------------------------------------
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak, which = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json)") if "hbm_gbs" in peaks else (6650.0, "fallback (B200_PROFILING.md)")
    traffic, traffic_note = None, "no capture committed for this kernel source"
    try:                          # dram bytes per launch from the committed ncu capture -- only if taken from THIS kernel source
        tr = json.load(open(os.path.join(ROOT, "profiles", "r2_slots_traffic.json")))
        if args.engine != 1 and tr.get("instances_per_launch") == B and tr.get("kernel_src_sha") == kernel_src_sha():
            traffic, traffic_note = tr["dram_bytes_per_launch"], f"ncu --set full capture {tr.get('capture', '')}, kernel source {tr['kernel_src_sha']}"
        else:
            traffic_note = f"committed capture is of kernel source {tr.get('kernel_src_sha')}, this build is {kernel_src_sha()}: stale, dropped"
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
        v_one, dt_one, _ = cpu_arm(48, 101, None)
        step_dev(0)
        torch.cuda.synchronize()
        parity = float(np.abs(u0.t().cpu().numpy()[:n_s] - u_cpu).max())
        cpu = {"value": v_cpu, "unit": "solves/s", "cores": cores, "kind": "port",
               "sample": f"first {n_s} instances of the timed batch ({dt_cpu:.1f} s) on {cores} cores; 48 instances on 1 core "
                         f"({dt_one:.1f} s); oracle port of solve_with_ltv (HiGHS + KKT polish); cvxpy/OSQP absent from the image",
               "value_1_core": v_one,
               "reference_shipped_log_ms_per_solve": 82.6,
               "reference_shipped_log_note": "logs/controls_20260208_014109.csv: CVXPY+OSQP on the authors' PC at the SMALLER "
                                             "N=6 / block 2 / M=3 shape, mean of 200 solves (= 12 solves/s)",
               "max_abs_u0_diff_vs_gpu": parity}

    secondary = None
    if world == 1 and not args.no_secondary:
        secondary = secondary_metrics(ctx, dev, stream, with_cpu=not args.no_cpu)

    line = {"metric": METRIC, "value": value, "unit": "solves/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "instances_per_gpu": B, "global_instances": gB,
                       "l2": f"inputs rotate over {R} distinct batches ({R}x{h2d_mat / 1e6:.0f} MB > 126 MB L2); kernel scratch "
                             f"{nat.load_library().b2c_mpc_scratch_bytes(C.byref(prm), B) / 1e6:.0f} MB",
                       "parallelism": f"batch-sharded x{world}, no data-path collective",
                       "host_binding_rank0": numa},
            "p99_batch_latency_ms": float(np.percentile(per_launch_ms, 99)),
            "p50_batch_latency_ms": float(np.percentile(per_launch_ms, 50)),
            "value_full_outputs": {"value": value_full, "unit": "solves/s", "ms_per_step": full_ms / Kf, "steps": Kf,
                                   "alg_bytes_per_solve": ALG_BYTES_FULL,
                                   "note": "control_sequence + predicted_states written, per-instance step_count supplied "
                                           "(what MPCController.solve_with_ltv always returns, mpc_controller.py:511-520)"},
            "solver": {"mean_newton_iters": iters_mean, "optimal_frac": ok_frac, "slack_active_frac": slack_frac,
                       "saturated_u0_frac": sat_frac},
            "roofline": {"bound": "fp64", "achieved": achieved_tf, "peak": fp64_peak, "unit": "TFLOP/s",
                         "frac": achieved_tf / fp64_peak, "traffic": traffic, "traffic_note": traffic_note,
                         "peak_source": "FMA microbenchmark in this run (b2c_measure_fma_peak; MEASURED_PEAKS.json has no FP64 entry)",
                         "kernel": kernel_name, "avg_launch_ms": avg_launch_s * 1e3,
                         "flop_per_solve_counted": flop_per_solve, "fp32_peak_tflops_measured": fp32_peak,
                         "why": "29 flop/B against a machine balance of 5 flop/B: FP64 is the binding roof, HBM is reported beside it"},
            "roofline_hbm": {"bound": "hbm", "achieved": achieved_gbs, "peak": hbm_peak, "unit": "GB/s",
                             "frac": achieved_gbs / hbm_peak, "peak_source": which,
                             "alg_bytes_per_solve": ALG_BYTES_PER_SOLVE},
            "e2e": {"value": gB * Ke / e2e_async_s, "unit": "solves/s", "h2d_bytes_per_step": h2d_table, "d2h_bytes_per_step": d2h,
                    "steps": Ke, "ms_per_step": 1e3 * e2e_async_s / Ke,
                    "api": "b2c_mpc_table_host_async + b2c_wait (C ABI, pinned host buffers, two calls in flight): shared "
                           "reference table + per-instance start row, segments gathered on the device",
                    "sync_value": gB * Ke / e2e_sync_s, "sync_ms_per_step": 1e3 * e2e_sync_s / Ke},
            "e2e_materialised": {"value": gB * Ke / e2e_mat_async_s, "unit": "solves/s", "h2d_bytes_per_step": h2d_mat,
                                 "d2h_bytes_per_step": d2h, "steps": Ke, "ms_per_step": 1e3 * e2e_mat_async_s / Ke,
                                 "api": "b2c_mpc_host_async + b2c_wait (C ABI, pinned host buffers, two calls in flight): every "
                                        "instance ships its own (N+1)-row segment -- bound by the host->device copy",
                                 "sync_value": gB * Ke / e2e_mat_s, "sync_ms_per_step": 1e3 * e2e_mat_s / Ke},
            "gpu_launches": int(launches),
            "clocks": sampler.summary(),
            "wall_s_timed_region": t_wall}
    if c5 is not None:
        line["config5"] = c5
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
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline legs")
    ap.add_argument("--no-secondary", action="store_true", help="skip the secondary (DARE / linearize / latency) figures")
    ap.add_argument("--no-config5", action="store_true", help="skip the config-5 closed-loop rollout leg")
    ap.add_argument("--config5-robots", type=int, default=1 << 20, help="robots per GPU of the config-5 leg")
    ap.add_argument("--config5-ticks", type=int, default=500)
    ap.add_argument("--engine", type=int, default=0, choices=[0, 1, 2],
                    help="K3 kernel: 0 auto (slot kernel for this shape), 1 general kernel, 2 slot kernel")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
