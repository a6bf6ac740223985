"""Oracle: the reference's closed-loop simulation loops restated on the CPU (numpy, fp64).  TEST INFRASTRUCTURE.

Follows /root/reference/run_simulation.py: run_lqr_simulation :58-96, run_mpc_simulation :225-293,
run_hybrid_simulation :500-566, with the risk supervisor of HC/controllers/risk_metrics.py:84-222 and the plant of
HC/models/differential_drive.py:138-172.  The controllers are the other oracle modules (lin_dare: scipy DARE, the
routine the reference itself calls; mpc_qp: the QP restatement, cvxpy/OSQP being absent from the image).

Pinned on the shipped run logs: LQR mode reproduces logs/states_20260207_234542.csv (G-LQR), MPC mode
logs/states_20260208_014109.csv (G-MPC) -- see tests/test_oracle_golden.py; the hybrid loop is pinned on the
switching pattern and states of logs/{states,controls}_20260208_003249.csv (tests/golden/g_hyb.npz)."""
import numpy as np

from . import lin_dare, mpc_qp


def distance_risk(px, py, obstacles, d_safe, d_trigger):
    """risk_metrics.py:84-132; obstacles (M,3) [x, y, radius]."""
    worst = 0.0
    for ox, oy, rad in obstacles:
        edge = np.sqrt((px - ox) ** 2 + (py - oy) ** 2) - rad
        if edge <= d_safe:
            r = 1.0
        elif edge >= d_trigger:
            r = 0.0
        else:
            r = 1.0 - (edge - d_safe) / (d_trigger - d_safe)
        worst = max(worst, r)
    return worst


def simulate(mode, table, k0, x0, obstacles, ticks, *, lqr_Q=(15.0, 15.0, 8.0), lqr_R=(0.1, 0.1), lqr_v_max=2.0,
             lqr_w_max=3.0, mpc_prm=None, dt=0.02, plant_v_max=2.0, plant_w_max=3.0, mpc_rate=5, min_dwell=10,
             risk=(0.3, 1.0, 0.6, 0.4, 0.2), step_count=0, mpc_solver=None):
    """One robot.  Returns dict(states (ticks+1,3), controls (ticks,2), controller (ticks,), risk (ticks,),
    errors (ticks,), switches, collision_states, step_count)."""
    table = np.asarray(table, float)
    rows = len(table)
    x = np.array(x0, float)
    obstacles = np.asarray(obstacles, float).reshape(-1, 3)
    d_safe_r, d_trig, alpha, beta, thr = risk
    alpha = alpha / (alpha + beta)
    solve = mpc_solver or (lambda x_, xr_, ur_, sc_: mpc_qp.solve_with_ltv(x_, xr_, ur_, obstacles, step_count=sc_, **mpc_prm))
    states = np.zeros((ticks + 1, 3)); states[0] = x
    controls = np.zeros((ticks, 2)); which = np.zeros(ticks, np.int8); rk = np.zeros(ticks); err = np.zeros(ticks)
    prev, since, switches, u = None, 0, 0, np.zeros(2)
    N = mpc_prm["N"] if mpc_prm else 0
    d_safe_mpc = mpc_prm["d_safe"] if mpc_prm else 0.0

    def collides(s):
        return any(np.sqrt((s[0] - o[0]) ** 2 + (s[1] - o[1]) ** 2) < o[2] + d_safe_mpc for o in obstacles)
    collisions = int(collides(x))
    for t in range(ticks):
        k = min(k0 + t, rows - 1)
        x_ref, u_ref = table[k, 1:4], table[k, 4:6]
        if mode == "lqr":
            use_mpc = False
        elif mode == "mpc":
            use_mpc = True
        else:
            rk[t] = alpha * distance_risk(x[0], x[1], obstacles, d_safe_r, d_trig)
            rec = rk[t] >= thr
            use_mpc = rec if (since >= min_dwell or prev is None) else (prev == 1)
            cur = 1 if use_mpc else 0
            if prev is not None and cur != prev:
                switches += 1; since = 0
            else:
                since += 1
            prev = cur
        which[t] = 1 if use_mpc else 0
        err[t] = np.hypot(x[0] - x_ref[0], x[1] - x_ref[1])
        if not use_mpc:
            K, _ = lin_dare.lqr_gain(u_ref[0], x_ref[2], lqr_Q, lqr_R, dt)
            u, _ = lin_dare.lqr_control(x, x_ref, u_ref, K, lqr_v_max, lqr_w_max)
        elif mode != "mpc" or t % mpc_rate == 0:
            xr, ur = lin_dare.segment(table, k0 + t, N + 1)
            sol = solve(x, xr, ur[:N], step_count)
            u = np.array(sol["optimal_control"], float)
            if sol["status"] == "optimal":
                step_count += 1
        controls[t] = u
        x = lin_dare.plant_step(x, u, dt, plant_v_max, plant_w_max)
        states[t + 1] = x
        collisions += int(collides(x))
    return dict(states=states, controls=controls, controller=which, risk=rk, errors=err, switches=switches,
                collision_states=collisions, step_count=step_count)
