"""CPU, world_size 2 over gloo: the batch-sharded path.  Each rank generates ITS slice of
the global batch, solves it (host build of the kernel code, tests/emul.py), and the pieces must be
bit-identical to the single-process solve of the same global indices; the statistics all_reduce must
equal the statistics of the whole batch."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import emul  # noqa: F401  (builds the emulation library once, before the ranks fork)
from conftest import ROOT, PKG

B, N = 256, 20


def _params():
    from hybrid_controller import _native as nat
    return nat.mpc_params(nat.MPC_LTV, N, 5, 1, [15, 15, 50], [.1, .1], [30, 30, 40], .3, 5000., 2., 3., .02)


def _rank_main(rank, world, port, outdir):
    for p in (ROOT, PKG, os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import emul as em
    from hybrid_controller import workloads as wl, dist_stats
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    start, count = wl.shard_range(B, rank, world)
    x0, xr, ur, ob = wl.config4_instances(B, N=N, seed=3, start=start, count=count)
    out = em.mpc(_params(), x0, xr, ur, ob)
    st = dist_stats.reduce_stats(dist_stats.solve_stats(out["status"], out["iters"], out["slack_used"], out["cost"]))
    np.savez(os.path.join(outdir, f"r{rank}.npz"), start=start, count=count, u0=out["u0"], cost=out["cost"],
             n=st["n"], n_slack=st["n_slack"], iters_sum=st["iters_sum"], iters_max=st["iters_max"],
             hist=st["iters_hist"], cost_sum=st["cost_sum"], cost_max=st["cost_max"])
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_equals_single(tmp_path, world):
    from hybrid_controller import workloads as wl, dist_stats
    lib = emul.lib()
    assert lib is not None
    port = 29500 + (os.getpid() % 2000) + world
    mp.start_processes(_rank_main, args=(world, port, str(tmp_path)), nprocs=world, join=True, start_method="fork")
    x0, xr, ur, ob = wl.config4_instances(B, N=N, seed=3)
    full = emul.mpc(_params(), x0, xr, ur, ob)
    ref = dist_stats.solve_stats(full["status"], full["iters"], full["slack_used"], full["cost"])
    covered = 0
    for r in range(world):
        d = np.load(os.path.join(tmp_path, f"r{r}.npz"))
        s, c = int(d["start"]), int(d["count"])
        assert np.array_equal(d["u0"], full["u0"][s:s + c]) and np.array_equal(d["cost"], full["cost"][s:s + c])
        covered += c
        assert int(d["n"]) == B and int(d["n_slack"]) == int(ref["n_slack"])
        assert int(d["iters_sum"]) == int(ref["iters_sum"]) and int(d["iters_max"]) == int(ref["iters_max"])
        assert np.array_equal(d["hist"], ref["iters_hist"])
        assert abs(float(d["cost_sum"]) - float(ref["cost_sum"])) <= 1e-9 * abs(float(ref["cost_sum"]))
        assert float(d["cost_max"]) == float(ref["cost_max"])
    assert covered == B


def test_shard_range_partitions():
    from hybrid_controller import workloads as wl
    for Bn in (65536, 1000, 7):
        for world in (1, 2, 3, 8):
            spans = [wl.shard_range(Bn, r, world) for r in range(world)]
            assert spans[0][0] == 0 and sum(c for _, c in spans) == Bn
            for (s0, c0), (s1, _) in zip(spans, spans[1:]):
                assert s0 + c0 == s1


# ---- config 5 (closed-loop rollouts): robots keyed by global index, rollout statistics all-reduced ----------------
R5, T5 = 96, 40


def _controllers5():
    from hybrid_controller.controllers import LQRController, MPCController, RiskMetrics
    from hybrid_controller import workloads as wl
    return dict(lqr=LQRController(**wl.SIM_LQR), mpc=MPCController(horizon=6, block_size=1, **wl.SIM_MPC),
                risk=RiskMetrics(d_safe=0.3, d_trigger=1.0, alpha=0.6, beta=0.4, threshold_low=0.2, threshold_medium=0.5))


def _rank_main5(rank, world, port, outdir):
    for p in (ROOT, PKG, os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import emul as em
    from hybrid_controller import workloads as wl, dist_stats
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    start, count = wl.shard_range(R5, rank, world)
    table = wl.figure8_table()
    k0, x0, ob = wl.config5_robots(count, first_index=start, seed=9, table=table)
    out = em.rollout("hybrid", table, k0, x0, ob, T5, **_controllers5())
    st = dist_stats.reduce_stats(dist_stats.rollout_stats(out, T5))
    np.savez(os.path.join(outdir, f"q{rank}.npz"), start=start, count=count, x=out["x"], mpc=out["mpc_solves"], **st)
    dist.barrier()
    dist.destroy_process_group()


def test_config5_sharded_rollout_equals_single(tmp_path):
    from hybrid_controller import workloads as wl, dist_stats
    world = 2
    port = 31500 + (os.getpid() % 2000)
    mp.start_processes(_rank_main5, args=(world, port, str(tmp_path)), nprocs=world, join=True, start_method="fork")
    table = wl.figure8_table()
    k0, x0, ob = wl.config5_robots(R5, first_index=0, seed=9, table=table)
    full = emul.rollout("hybrid", table, k0, x0, ob, T5, **_controllers5())
    ref = dist_stats.rollout_stats(full, T5)
    assert ref["mpc_solves"] > 0 and ref["lqr_steps"] > 0
    for r in range(world):
        d = np.load(os.path.join(tmp_path, f"q{r}.npz"))
        s, c = int(d["start"]), int(d["count"])
        assert np.array_equal(d["x"], full["x"][s:s + c]) and np.array_equal(d["mpc"], full["mpc_solves"][s:s + c])
        for k in ("robots", "robot_steps", "lqr_steps", "mpc_solves", "switches", "collision_states", "mpc_fallbacks"):
            assert int(d[k]) == int(ref[k]), k
        assert abs(float(d["err_mean_sum"]) - float(ref["err_mean_sum"])) <= 1e-9 * float(ref["err_mean_sum"])
        assert float(d["err_max"]) == float(ref["err_max"])
