"""Batch statistics and their (optional) reduction across ranks (the only collective).

Instances are independent, so nothing crosses GPUs on the data path; at the end of a run each rank
holds a <= 1 KB statistics vector (counts, iteration histogram, sums, maxima) that is combined with
one all_reduce(SUM) and one all_reduce(MAX) -- NCCL over NVLink on GPUs, gloo in the CPU tests.
"""
import numpy as np

HIST_BINS = 32


def solve_stats(status, iters, slack_used, cost=None):
    """Per-shard statistics of a batch of MPC solves -> dict of numpy scalars / small arrays."""
    status = np.asarray(status)
    iters = np.asarray(iters)
    ok = status == 0
    s = {"n": np.int64(status.size), "n_optimal": np.int64(ok.sum()),
         "n_fallback": np.int64((~ok).sum()), "n_slack": np.int64(np.asarray(slack_used)[ok].sum()),
         "iters_sum": np.int64(iters.sum()), "iters_max": np.int64(iters.max() if iters.size else 0),
         "iters_hist": np.bincount(np.minimum(iters, HIST_BINS - 1), minlength=HIST_BINS).astype(np.int64)}
    if cost is not None:
        c = np.asarray(cost)[ok]
        s["cost_sum"] = np.float64(c.sum())
        s["cost_max"] = np.float64(c.max() if c.size else -np.inf)
    return s


def rollout_stats(out, ticks):
    """Per-shard statistics of a closed-loop rollout (simulate_batch / simulate_generated_shard result) -> dict of
    numpy scalars: the <= 1 KB vector that is all-reduced over the ranks at the end of a config-5 run."""
    R = len(out["lqr_steps"])
    return {"robots": np.int64(R), "robot_steps": np.int64(R) * np.int64(ticks),
            "lqr_steps": np.int64(out["lqr_steps"].sum(dtype=np.int64)), "mpc_solves": np.int64(out["mpc_solves"].sum(dtype=np.int64)),
            "switches": np.int64(out["switches"].sum(dtype=np.int64)),
            "collision_states": np.int64(out["collision_states"].sum(dtype=np.int64)),
            "mpc_fallbacks": np.int64(out["mpc_fallbacks"].sum(dtype=np.int64)),
            "err_mean_sum": np.float64(out["mean_error"].sum()), "err_max": np.float64(out["max_error"].max())}


_MAX_KEYS = ("iters_max", "cost_max", "err_max")


def reduce_stats(stats, device=None):
    """all_reduce the statistics over the default process group (no-op when not initialised)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return stats
    sum_keys = [k for k in stats if k not in _MAX_KEYS]
    max_keys = [k for k in stats if k in _MAX_KEYS]
    out = dict(stats)
    for keys, op in ((sum_keys, dist.ReduceOp.SUM), (max_keys, dist.ReduceOp.MAX)):
        if not keys:
            continue
        flat = np.concatenate([np.atleast_1d(np.asarray(stats[k], dtype=np.float64)) for k in keys])
        t = torch.from_numpy(flat)
        if device is not None:
            t = t.to(device)
        dist.all_reduce(t, op=op)
        flat = t.cpu().numpy()
        off = 0
        for k in keys:
            n = np.atleast_1d(stats[k]).size
            v = flat[off:off + n].astype(np.asarray(stats[k]).dtype)
            out[k] = v if np.ndim(stats[k]) else v[0]
            off += n
    return out
