"""Batched closed-loop rollouts on the device: the reference's three simulation loops (run_simulation.py
run_lqr_simulation :58-96, run_mpc_simulation :225-280, run_hybrid_simulation :500-570) for R robots at once.

    out = simulate_batch("hybrid", table, k0, x0, obstacles, ticks, lqr=LQRController(...), mpc=MPCController(...),
                         risk=RiskMetrics(...))

Every robot follows the shared reference `table` (ReferenceTrajectoryGenerator.generate()) from its own start row
k0[r] and initial state x0[r], with its own obstacle set.  The per-robot state of the reference's objects (hysteresis
counters, MPCController._step_count) lives on the device for the duration of the call."""
import ctypes as C
from typing import Any, Dict, Optional

import numpy as np

from . import _native as nat

MODES = {"lqr": nat.ROLL_LQR, "mpc": nat.ROLL_MPC, "hybrid": nat.ROLL_HYBRID}


def rollout_params(mode, ticks, table_rows, lqr=None, mpc=None, risk=None, n_obstacles=0, dt=0.02,
                   plant_v_max=2.0, plant_w_max=3.0, mpc_rate=5, min_dwell=10):
    p = nat.RolloutParams()
    p.mode, p.ticks, p.mpc_rate, p.min_dwell, p.table_rows = MODES[mode], int(ticks), int(mpc_rate), int(min_dwell), int(table_rows)
    p.dt, p.plant_v_max, p.plant_w_max = float(dt), float(plant_v_max), float(plant_w_max)
    if lqr is not None:
        p.lqr_Qd[:] = np.diag(lqr.Q).tolist()
        p.lqr_Rd[:] = np.diag(lqr.R).tolist()
        p.lqr_v_max, p.lqr_w_max = float(lqr.v_max), float(lqr.omega_max)
    else:
        p.lqr_Qd[:] = [1.0, 1.0, 1.0]
        p.lqr_Rd[:] = [1.0, 1.0]
    if risk is not None:                             # RiskMetrics already holds the normalised weights (:76-79)
        p.risk_d_safe, p.risk_d_trigger = float(risk.d_safe), float(risk.d_trigger)
        p.risk_alpha, p.risk_beta, p.risk_threshold_low = float(risk.alpha), float(risk.beta), float(risk.threshold_low)
    else:
        p.risk_alpha, p.risk_beta = 1.0, 0.0
    if mpc is not None:
        p.mpc = mpc._params(nat.MPC_LTV, int(n_obstacles), soft=True)
    else:
        p.mpc.M = int(n_obstacles)
        p.mpc.N, p.mpc.kind = 1, nat.MPC_LTV
    return p


def simulate_batch(mode: str, table, k0, x0, obstacles, ticks: int, lqr=None, mpc=None, risk=None,
                   dt: float = 0.02, plant_v_max: float = 2.0, plant_w_max: float = 3.0, mpc_rate: int = 5,
                   min_dwell: int = 10, mpc_step_count=None, log: bool = False, ctx=None) -> Dict[str, Any]:
    """R robots x `ticks` plant steps.  table (rows,6); k0 (R,) int; x0 (R,3); obstacles (R,M,3) or (M,3) [x,y,radius].

    Returns final states `x` (R,3), counters `lqr_steps`, `mpc_solves`, `switches`, `collision_states`,
    `mpc_fallbacks` (R,), tracking statistics `mean_error`, `max_error`, `final_error` (R,) (position norm, as the
    reference prints them), the updated `mpc_step_count`, and with log=True the per-tick `states` (R,ticks+1,3),
    `controls` (R,ticks,2), `risk` (R,ticks), `controller` (R,ticks) [0 LQR / 1 MPC]."""
    ctx = ctx or nat.default_context()
    table = np.ascontiguousarray(table, np.float64)
    x0 = np.ascontiguousarray(x0, np.float64).reshape(-1, 3)
    R = x0.shape[0]
    k0 = np.ascontiguousarray(np.broadcast_to(np.asarray(k0, np.int32), (R,)))
    obstacles = np.asarray(obstacles, np.float64)
    if obstacles.ndim == 2:
        obstacles = np.broadcast_to(obstacles[None], (R,) + obstacles.shape)
    M = obstacles.shape[1] if obstacles.size else 0
    p = rollout_params(mode, ticks, table.shape[0], lqr, mpc, risk, M, dt, plant_v_max, plant_w_max, mpc_rate, min_dwell)
    x = np.array(x0.T, dtype=np.float64, order="C", copy=True)       # [3][R]; a copy even when R == 1 (the caller's x0 stays untouched)
    ob = np.ascontiguousarray(obstacles.reshape(R, 3 * M).T) if M else np.zeros((1, R))
    sc = None if mpc_step_count is None else np.ascontiguousarray(mpc_step_count, np.int32).copy()
    counters, stats = np.zeros((5, R), np.int32), np.zeros((3, R))
    lx = np.empty((ticks + 1, 3, R)) if log else None
    lu = np.empty((ticks, 2, R)) if log else None
    lr = np.empty((ticks, R)) if log else None
    lc = np.empty((ticks, R), np.int8) if log else None
    ctx.call("b2c_rollout_host", R, C.byref(p), nat._ptr(table), nat._ptr(k0), nat._ptr(x), nat._ptr(ob), nat._ptr(sc),
             nat._ptr(counters), nat._ptr(stats), nat._ptr(lx), nat._ptr(lu), nat._ptr(lr), nat._ptr(lc))
    out = dict(x=x.T.copy(), lqr_steps=counters[0], mpc_solves=counters[1], switches=counters[2],
               collision_states=counters[3], mpc_fallbacks=counters[4],
               mean_error=stats[0] / max(ticks, 1), max_error=stats[1], final_error=stats[2], mpc_step_count=sc)
    if log:
        out.update(states=lx.transpose(2, 0, 1).copy(), controls=lu.transpose(2, 0, 1).copy(), risk=lr.T.copy(),
                   controller=lc.T.copy())
    return out


def simulate_generated_shard(R: int, first_index: int, seed: int, table, ticks: int, mode: str = "hybrid", lqr=None,
                             mpc=None, risk=None, obstacles=None, k0_max: int = 500, pos_jitter: float = 0.1,
                             th_jitter: float = 0.2, obs_jitter: float = 0.2, ctx=None, device=None, return_inputs: bool = False,
                             **loop_kw) -> Dict[str, Any]:
    """One shard of BASELINE config 5, entirely on the device: robots [first_index, first_index + R) are GENERATED there
    (b2c_rollout_generate_dev: counter-based stream keyed by (seed, global robot index), so the shard split never
    changes a robot) and rolled out by b2c_rollout_dev; only the per-robot counters / statistics come back.
    Device memory comes from torch (imported here, not at package import).  Returns the simulate_batch dict without
    logs, plus `elapsed_s` (CUDA events around generation + rollout on the context's stream)."""
    import torch
    from . import workloads as wl
    dev = torch.device("cuda", torch.cuda.current_device() if device is None else device)
    ctx = ctx or nat.default_context(dev.index)
    table = np.ascontiguousarray(table, np.float64)
    base = np.ascontiguousarray(wl.DEFAULT_OBSTACLES if obstacles is None else obstacles, np.float64).reshape(-1, 3)
    M = base.shape[0]
    p = rollout_params(mode, ticks, table.shape[0], lqr, mpc, risk, M, **loop_kw)
    d_tab = torch.from_numpy(table).to(dev)
    d_k0 = torch.empty(R, dtype=torch.int32, device=dev)
    d_x = torch.empty((3, R), dtype=torch.float64, device=dev)
    d_ob = torch.empty((max(3 * M, 1), R), dtype=torch.float64, device=dev)
    cnt = torch.empty((5, R), dtype=torch.int32, device=dev)
    sts = torch.empty((3, R), dtype=torch.float64, device=dev)
    torch.cuda.synchronize(dev)
    stream = torch.cuda.ExternalStream(ctx.stream_handle, device=dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    ctx.call("b2c_rollout_generate_dev", R, int(first_index), int(seed), d_tab.data_ptr(), table.shape[0], int(k0_max),
             nat._ptr(base), M, float(pos_jitter), float(th_jitter), float(obs_jitter), d_k0.data_ptr(), d_x.data_ptr(),
             d_ob.data_ptr())
    inputs = (d_k0.cpu().numpy(), d_x.t().cpu().numpy().copy(), d_ob[:3 * M].t().cpu().numpy().reshape(R, M, 3).copy()) \
        if return_inputs else None
    ctx.call("b2c_rollout_dev", R, C.byref(p), d_tab.data_ptr(), d_k0.data_ptr(), d_x.data_ptr(), d_ob.data_ptr(), None,
             cnt.data_ptr(), sts.data_ptr(), None, None, None, None)
    e1.record(stream)
    torch.cuda.synchronize(dev)
    c, st = cnt.cpu().numpy(), sts.cpu().numpy()
    out = dict(x=d_x.t().cpu().numpy().copy(), lqr_steps=c[0], mpc_solves=c[1], switches=c[2], collision_states=c[3],
               mpc_fallbacks=c[4], mean_error=st[0] / max(ticks, 1), max_error=st[1], final_error=st[2],
               elapsed_s=e0.elapsed_time(e1) * 1e-3)
    if return_inputs:
        out["inputs"] = inputs
    return out


def export_robot_csv(out: Dict[str, Any], r: int, table, k0: int, log_dir: str, stamp: str, mode: str = "hybrid",
                     wrap_error: Optional[bool] = None):
    """Write robot r of a simulate_batch(log=True) result in the reference's run-log schema
    (HC/logging/simulation_logger.py:174-186 states, :221-227 controls, csv.DictWriter as :402-452), so that one robot
    of a batch can be diffed against a reference run with the tooling used on its shipped logs.

    states_<stamp>.csv  : timestep,px,py,theta,px_ref,py_ref,theta_ref,error_px,error_py,error_theta,error_norm
    controls_<stamp>.csv: timestep,v,omega,controller,solve_time_ms
    The heading error is wrapped in the lqr / mpc loops and left raw in the hybrid loop (run_simulation.py:562, Q12)."""
    import csv
    import os
    table = np.asarray(table, float)
    states, controls, which = out["states"][r], out["controls"][r], out["controller"][r]
    wrap = (mode != "hybrid") if wrap_error is None else wrap_error
    os.makedirs(log_dir, exist_ok=True)
    ps, pc = os.path.join(log_dir, f"states_{stamp}.csv"), os.path.join(log_dir, f"controls_{stamp}.csv")
    with open(ps, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["timestep", "px", "py", "theta", "px_ref", "py_ref", "theta_ref", "error_px", "error_py",
                    "error_theta", "error_norm"])
        for t in range(len(controls)):
            ref = table[min(k0 + t, len(table) - 1), 1:4]
            e = states[t] - ref
            if wrap:
                while e[2] > np.pi:
                    e[2] -= 2 * np.pi
                while e[2] < -np.pi:
                    e[2] += 2 * np.pi
            w.writerow([t] + [float(v) for v in states[t]] + [float(v) for v in ref] + [float(v) for v in e]
                       + [float(np.linalg.norm(e))])
    with open(pc, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["timestep", "v", "omega", "controller", "solve_time_ms"])
        for t in range(len(controls)):
            w.writerow([t, float(controls[t, 0]), float(controls[t, 1]), "MPC" if which[t] else "LQR", ""])
    return ps, pc
