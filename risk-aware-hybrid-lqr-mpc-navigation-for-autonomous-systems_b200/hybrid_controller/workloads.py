"""Synthetic batched workloads of BASELINE.json / SURVEY.md 8(d) (host-side numpy generators).

  config 3: B linearization points along the figure-8 (+ a seeded stress set) for K1 / K2
  config 4: B independent diff-drive MPC instances, N = 20, 5 'dense' obstacles, for K3
Sharding rule (8e): instances are drawn for the GLOBAL index range and sliced, so the instance a
global index names is the same for every world size.
"""
import numpy as np

from .trajectory.reference_generator import ReferenceTrajectoryGenerator

# run_simulation.py:198-204 ("dense" scenario) and :215-219 ("default")
DENSE_OBSTACLES = np.array([[1.0, 0.5, 0.2], [-0.5, -1.0, 0.25], [1.5, -0.3, 0.15],
                            [-1.5, 0.5, 0.2], [0.0, 0.8, 0.15]])
DEFAULT_OBSTACLES = DENSE_OBSTACLES[:3].copy()

# run_simulation.py:164-176 (MPC) and :54 (LQR)
SIM_MPC = dict(Q_diag=[15.0, 15.0, 50.0], R_diag=[0.1, 0.1], P_diag=[30.0, 30.0, 40.0], d_safe=0.3,
               slack_penalty=5000.0, dt=0.02, v_max=2.0, omega_max=3.0)
SIM_LQR = dict(Q_diag=[15.0, 15.0, 8.0], R_diag=[0.1, 0.1], dt=0.02, v_max=2.0, omega_max=3.0)

_traj_cache = {}


def figure8_table(duration=20.0, dt=0.02, A=2.0, a=0.5):
    key = (duration, dt, A, a)
    if key not in _traj_cache:
        _traj_cache[key] = ReferenceTrajectoryGenerator(A=A, a=a, dt=dt).generate(duration)
    return _traj_cache[key]


def config3_points(B, stress=False, seed=0, A=2.0, a=0.5):
    """(v, theta) operating points: one figure-8 period sampled at B phases, or the stress set."""
    if stress:
        rng = np.random.default_rng(seed)
        return rng.uniform(0.01, 2.0, B), rng.uniform(-np.pi, np.pi, B)
    t = np.arange(B) * ((2 * np.pi / a) / B)
    c, s = np.cos(a * t), np.sin(a * t)
    dpx, dpy = a * A * c, a * A * (c**2 - s**2)
    return np.sqrt(dpx**2 + dpy**2), np.arctan2(dpy, dpx)


def config4_instances(B, N=20, seed=0, start=0, count=None, obstacles=DENSE_OBSTACLES,
                      pos_jitter=0.1, th_jitter=0.2, obs_jitter=0.2, return_k0=False):
    """MPC instances [start, start+count) of a global batch of B (instance-major arrays).

    Per instance: k0 ~ U{0..979}; (x_refs, u_refs) = segment(k0, N+1) of generate(20.0);
    x0 = x_refs[0] + U(+-.1, +-.1, +-.2); the obstacle set shifted by U(+-.2)^2 per obstacle.
    Returns x0 (n,3), x_refs (n,N+1,3), u_refs (n,N,2), obs (n,M,3) [, k0 (n,) int32: x_refs / u_refs are
    get_trajectory_segment(k0, N+1) of figure8_table()]."""
    count = B - start if count is None else count
    table = figure8_table()
    rng = np.random.default_rng(seed)
    k0 = rng.integers(0, 980, size=B)
    dx0 = np.stack([rng.uniform(-pos_jitter, pos_jitter, B), rng.uniform(-pos_jitter, pos_jitter, B),
                    rng.uniform(-th_jitter, th_jitter, B)], axis=1)
    M = len(obstacles)
    dob = rng.uniform(-obs_jitter, obs_jitter, (B, M, 2))
    sl = slice(start, start + count)
    idx = np.minimum(k0[sl, None] + np.arange(N + 1)[None, :], len(table) - 1)
    x_refs = table[idx][:, :, 1:4].copy()
    u_refs = table[idx][:, :N, 4:6].copy()
    x0 = x_refs[:, 0, :] + dx0[sl]
    obs = np.broadcast_to(np.asarray(obstacles, float)[None], (count, M, 3)).copy()
    obs[:, :, :2] += dob[sl]
    if return_k0:                 # the table form of the entry points takes (table, k0) instead of the segments
        return x0, x_refs, u_refs, obs, k0[sl].astype(np.int32)
    return x0, x_refs, u_refs, obs


_GOLDEN = np.uint64(0x9E3779B97F4A7C15)


def _u01(state, j):
    """draw j of the counter-based stream (numpy mirror of gen_u01 in csrc/rollout.cu; uint64 arithmetic wraps)"""
    with np.errstate(over="ignore"):
        z = state + np.uint64(j + 1) * _GOLDEN
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return (z >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def config5_robots(R, first_index=0, seed=0, table=None, k0_max=500, obstacles=DEFAULT_OBSTACLES,
                   pos_jitter=0.1, th_jitter=0.2, obs_jitter=0.2):
    """Robots [first_index, first_index + R) of BASELINE config 5, keyed by GLOBAL robot index: the numpy
    mirror of b2c_rollout_generate_dev, bit-identical to it.  Returns k0 (R,) int32, x0 (R,3), obs (R,M,3)."""
    table = figure8_table() if table is None else np.asarray(table, float)
    obstacles = np.asarray(obstacles, float).reshape(-1, 3)
    with np.errstate(over="ignore"):
        state = np.uint64(seed) ^ ((np.arange(R, dtype=np.uint64) + np.uint64(first_index)) * _GOLDEN)
    k0 = np.minimum((_u01(state, 0) * float(k0_max)).astype(np.int32), k0_max - 1)
    rows = table[np.minimum(k0, len(table) - 1)]
    jit = (pos_jitter, pos_jitter, th_jitter)
    x0 = np.stack([rows[:, 1 + c] + (2.0 * _u01(state, 1 + c) - 1.0) * jit[c] for c in range(3)], axis=1)
    M = len(obstacles)
    obs = np.empty((R, M, 3))
    for m in range(M):
        obs[:, m, 0] = obstacles[m, 0] + (2.0 * _u01(state, 4 + 2 * m) - 1.0) * obs_jitter
        obs[:, m, 1] = obstacles[m, 1] + (2.0 * _u01(state, 5 + 2 * m) - 1.0) * obs_jitter
        obs[:, m, 2] = obstacles[m, 2]
    return k0, x0, obs


def shard_range(B, rank, world):
    """Contiguous block split of the global index range."""
    per = B // world
    start = rank * per
    return start, (B - start if rank == world - 1 else per)
