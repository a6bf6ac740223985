"""Test-only: run the per-instance kernel code (mpc_solve_one / dare_solve_one, compiled for the host
with -DB2C_HOST_EMUL into tests/_emul/libb2c_emul.so) on the CPU.  Not part of the product."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "risk-aware-hybrid-lqr-mpc-navigation-for-autonomous-systems_b200", "csrc")
SO = os.path.join(ROOT, "tests", "_emul", "libb2c_emul.so")
_lib = None


def lib():
    global _lib
    if _lib is None:
        srcs = [os.path.join(CSRC, f) for f in ("mpc.cu", "mpc_slots.cu", "mpc_args.cuh", "lin_dare.cu", "api.cu", "b2c_common.cuh", "rollout.cu")]
        if not os.path.exists(SO) or any(os.path.getmtime(s) > os.path.getmtime(SO) for s in srcs):
            subprocess.check_call(["make", "-C", CSRC, "emul"], stdout=subprocess.DEVNULL)
        _lib = C.CDLL(SO)
    return _lib


def _soa(a, B):
    return np.ascontiguousarray(np.asarray(a, float).reshape(B, -1).T)


def mpc(params, x0, xref, uref, obs, step_count=None, slots=False):
    from hybrid_controller import _native as nat
    N, M = params.N, params.M
    x0 = np.asarray(x0, float).reshape(-1, 3)
    B = x0.shape[0]
    sx0, sxr, sur = _soa(x0, B), _soa(xref, B), _soa(uref, B)
    sob = _soa(obs, B) if M else np.zeros((1, B))
    sc = None if step_count is None else np.ascontiguousarray(step_count, np.int32).copy()
    u0, cost = np.empty((2, B)), np.empty(B)
    st, it, sl = (np.empty(B, np.int32) for _ in range(3))
    useq, xpred = np.empty((2 * N, B)), np.empty((3 * (N + 1), B))
    p = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)
    f = lib().b2c_emul_mpc_slots if slots else lib().b2c_emul_mpc
    f.argtypes = [C.c_int64, C.POINTER(nat.MpcParams)] + [C.c_void_p] * 12
    rc = f(B, C.byref(params), p(sx0), p(sxr), p(sur), p(sob), p(sc), p(u0), p(cost), p(st), p(it),
           p(sl), p(useq), p(xpred))
    assert rc == 0, rc
    return dict(u0=u0.T.copy(), cost=cost, status=st, iters=it, slack_used=sl,
                useq=useq.T.reshape(B, N, 2).copy(), xpred=xpred.T.reshape(B, N + 1, 3).copy(),
                step_count=sc)


def mpc_slots_ex(params, x0, obs, xref=None, uref=None, table=None, k0=None, step_count=None):
    """The slot state machine with the layouts the host entry points use: numpy instance-major inputs / outputs
    (in_aos = out_aos = 1) and, with `table` + `k0`, the table form (segments gathered from the shared table)."""
    from hybrid_controller import _native as nat
    N, M = params.N, params.M
    x0 = np.ascontiguousarray(np.asarray(x0, float).reshape(-1, 3))
    B = x0.shape[0]
    obs = np.ascontiguousarray(np.asarray(obs, float).reshape(B, -1)) if M else np.zeros((B, 1))
    xr = None if xref is None else np.ascontiguousarray(np.asarray(xref, float).reshape(B, -1))
    ur = None if uref is None else np.ascontiguousarray(np.asarray(uref, float).reshape(B, -1))
    tab = None if table is None else np.ascontiguousarray(table, float)
    kk = None if k0 is None else np.ascontiguousarray(k0, np.int32)
    sc = None if step_count is None else np.ascontiguousarray(step_count, np.int32).copy()
    u0, cost = np.empty((B, 2)), np.empty(B)
    st, it, sl = (np.empty(B, np.int32) for _ in range(3))
    useq, xpred = np.empty((B, N, 2)), np.empty((B, N + 1, 3))
    p = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)
    f = lib().b2c_emul_mpc_slots_ex
    f.argtypes = [C.c_int64, C.POINTER(nat.MpcParams), C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p] + [C.c_void_p] * 12
    rc = f(B, C.byref(params), 1, 1, p(tab), 0 if tab is None else tab.shape[0], p(kk), p(x0), p(xr), p(ur), p(obs),
           p(sc), p(u0), p(cost), p(st), p(it), p(sl), p(useq), p(xpred))
    assert rc == 0, rc
    return dict(u0=u0, cost=cost, status=st, iters=it, slack_used=sl, useq=useq, xpred=xpred, step_count=sc)


def dare(v, th, dt, Qd, Rd, method=1, guard=1):
    v, th = np.ascontiguousarray(v, float).ravel(), np.ascontiguousarray(th, float).ravel()
    B = v.size
    K, P = np.empty((6, B)), np.empty((6, B))
    it, st = np.empty(B, np.int32), np.empty(B, np.int32)
    f = lib().b2c_emul_dare
    f.argtypes = [C.c_int64, C.c_int, C.c_int, C.c_double, C.c_double * 3, C.c_double * 2] + [C.c_void_p] * 6
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    f(B, method, guard, dt, (C.c_double * 3)(*map(float, Qd)), (C.c_double * 2)(*map(float, Rd)),
      p(v), p(th), p(K), p(P), p(it), p(st))
    Pf = np.empty((B, 3, 3))
    for (r, c), e in zip(((0, 0), (0, 1), (0, 2), (1, 1), (1, 2), (2, 2)), range(6)):
        Pf[:, r, c] = P[e]; Pf[:, c, r] = P[e]
    return K.T.reshape(B, 2, 3).copy(), Pf, it, st


def linearize(v, th, dt, method=1, guard=0):
    v, th = np.ascontiguousarray(v, float).ravel(), np.ascontiguousarray(th, float).ravel()
    B = v.size
    A, Bd = np.empty((9, B)), np.empty((6, B))
    f = lib().b2c_emul_linearize
    f.argtypes = [C.c_int64, C.c_int, C.c_int, C.c_double] + [C.c_void_p] * 4
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    f(B, method, guard, dt, p(v), p(th), p(A), p(Bd))
    return A.T.reshape(B, 3, 3).copy(), Bd.T.reshape(B, 3, 2).copy()


def rollout(mode, table, k0, x0, obstacles, ticks, lqr=None, mpc=None, risk=None, mpc_step_count=None, **kw):
    """Host emulation of b2c_rollout_dev (same per-robot code, emulated K2/K3); same return dict as
    hybrid_controller.rollout.simulate_batch(log=True)."""
    from hybrid_controller import _native as nat, rollout as ro
    table = np.ascontiguousarray(table, np.float64)
    x0 = np.ascontiguousarray(x0, np.float64).reshape(-1, 3)
    R = x0.shape[0]
    k0 = np.ascontiguousarray(np.broadcast_to(np.asarray(k0, np.int32), (R,)))
    obstacles = np.asarray(obstacles, np.float64)
    if obstacles.ndim == 2:
        obstacles = np.broadcast_to(obstacles[None], (R,) + obstacles.shape)
    M = obstacles.shape[1] if obstacles.size else 0
    p = ro.rollout_params(mode, ticks, table.shape[0], lqr, mpc, risk, M, **kw)
    x = np.ascontiguousarray(x0.T)
    ob = np.ascontiguousarray(obstacles.reshape(R, 3 * M).T) if M else np.zeros((1, R))
    sc = None if mpc_step_count is None else np.ascontiguousarray(mpc_step_count, np.int32).copy()
    counters, stats = np.zeros((5, R), np.int32), np.zeros((3, R))
    lx, lu = np.empty((ticks + 1, 3, R)), np.empty((ticks, 2, R))
    lr, lc = np.empty((ticks, R)), np.empty((ticks, R), np.int8)
    f = lib().b2c_emul_rollout
    f.argtypes = [C.c_int64, C.POINTER(nat.RolloutParams)] + [C.c_void_p] * 11
    ptr = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)
    rc = f(R, C.byref(p), ptr(table), ptr(k0), ptr(x), ptr(ob), ptr(sc), ptr(counters), ptr(stats), ptr(lx), ptr(lu),
           ptr(lr), ptr(lc))
    assert rc == 0, rc
    return dict(x=x.T.copy(), lqr_steps=counters[0], mpc_solves=counters[1], switches=counters[2],
                collision_states=counters[3], mpc_fallbacks=counters[4], mean_error=stats[0] / max(ticks, 1),
                max_error=stats[1], final_error=stats[2], mpc_step_count=sc,
                states=lx.transpose(2, 0, 1).copy(), controls=lu.transpose(2, 0, 1).copy(), risk=lr.T.copy(),
                controller=lc.T.copy())
