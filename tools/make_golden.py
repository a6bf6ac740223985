#!/usr/bin/env python3
"""Generate tests/golden/*.npz from the UNMODIFIED reference at /root/reference.

Run in the build container only (the GPU box has no /root/reference):
    python tools/make_golden.py
Inputs : the reference's own Python (imported, cvxpy/matplotlib stubbed because they are absent
         and only needed by code we do not call) and the run logs it ships under logs/.
Outputs: tests/golden/g_lqr.npz   G-LQR  run log 20260207_234542 (999 ticks)
         tests/golden/g_mpc.npz   G-MPC  run log 20260208_014109 (999 ticks, 200 solves)
         tests/golden/ref_py.npz  outputs of the reference's Linearizer / LQRController /
                                  ReferenceTrajectoryGenerator / DifferentialDriveRobot on seeded inputs
"""
import csv
import os
import re
import sys
import types

import numpy as np

REF = os.environ.get("REF_ROOT", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")


def import_reference():
    for name in ("cvxpy", "matplotlib", "matplotlib.pyplot", "matplotlib.patches",
                 "matplotlib.collections"):
        if name not in sys.modules:
            m = types.ModuleType(name)
            m.__getattr__ = lambda attr: (_ for _ in ()).throw(AttributeError(attr))
            sys.modules[name] = m
    sys.modules["matplotlib"].use = lambda *a, **k: None
    sys.path.insert(0, os.path.join(REF, "src", "hybrid_controller"))
    from hybrid_controller.models.linearization import Linearizer
    from hybrid_controller.controllers.lqr_controller import LQRController
    from hybrid_controller.trajectory.reference_generator import ReferenceTrajectoryGenerator
    from hybrid_controller.models.differential_drive import DifferentialDriveRobot
    return Linearizer, LQRController, ReferenceTrajectoryGenerator, DifferentialDriveRobot


def read_log(stamp):
    st = np.array([[float(v) for v in r[:10]] for r in list(csv.reader(
        open(f"{REF}/logs/states_{stamp}.csv")))[1:]])
    rows = list(csv.reader(open(f"{REF}/logs/controls_{stamp}.csv")))[1:]
    ct = np.array([[float(r[0]), float(r[1]), float(r[2]), float(r[4]) if r[4] else np.nan]
                   for r in rows])
    slack = []
    for line in open(f"{REF}/logs/simulation_{stamp}.log"):
        m = re.search(r"k=(\d+) \| Constraint: slack_activated", line)
        if m:
            slack.append(int(m.group(1)))
    return st, ct, np.array(slack, dtype=np.int64), open(f"{REF}/logs/simulation_{stamp}.log").read()


def main():
    os.makedirs(OUT, exist_ok=True)
    Linearizer, LQRController, Gen, Robot = import_reference()

    st, ct, _, _ = read_log("20260207_234542")
    np.savez_compressed(os.path.join(OUT, "g_lqr.npz"), states=st[:, 1:4], refs=st[:, 4:7],
                        errors=st[:, 7:10], controls=ct[:, 1:3],
                        source="logs/states_20260207_234542.csv + controls_20260207_234542.csv")
    st, ct, slack, logtxt = read_log("20260208_014109")
    print("G-MPC slack ticks parsed:", slack.tolist())
    np.savez_compressed(os.path.join(OUT, "g_mpc.npz"), states=st[:, 1:4], refs=st[:, 4:7],
                        errors=st[:, 7:10], controls=ct[:, 1:3], solve_time_ms=ct[:, 3],
                        slack_ticks=slack,
                        source="logs/{states,controls,simulation}_20260208_014109.*")

    rng = np.random.default_rng(20260208)
    n = 256
    v = np.concatenate([rng.uniform(-2, 2, n - 6), [0.0, 1e-7, -1e-7, 0.01, 1.0, 1.0]])
    th = np.concatenate([rng.uniform(-np.pi, np.pi, n - 6), [0.3, -2.0, 1.0, np.pi, 0.0, np.pi / 4]])
    lin = Linearizer(dt=0.02)
    keys = ["jac", "euler", "exact", "explicit"]
    res = {k + "_A": np.zeros((n, 3, 3)) for k in keys}
    res.update({k + "_B": np.zeros((n, 3, 2)) for k in keys})
    for i in range(n):
        res["jac_A"][i], res["jac_B"][i] = lin.get_jacobians(v[i], th[i])
        res["euler_A"][i], res["euler_B"][i] = lin.get_discrete_model(v[i], th[i], "euler")
        res["exact_A"][i], res["exact_B"][i] = lin.get_discrete_model(v[i], th[i], "exact")
        res["explicit_A"][i], res["explicit_B"][i] = lin.get_discrete_model_explicit(v[i], th[i])
    # rollouts (linearization.py:227-280) and condensed matrices (:282-321)
    ctr = rng.uniform(-1, 1, (8, 2)); vr = rng.uniform(0.2, 1.5, 8); tr = rng.uniform(-3, 3, 8)
    x0 = np.array([0.1, -0.2, 0.3])
    res["roll_ctrl"], res["roll_v"], res["roll_th"], res["roll_x0"] = ctr, vr, tr, x0
    res["roll_ltv"] = lin.predict_trajectory(x0, ctr, vr, tr)
    res["roll_lti"] = lin.predict_horizon(x0, ctr, vr[0], tr[0])
    Phi, Gam = Linearizer.build_prediction_matrices(res["explicit_A"][5], res["explicit_B"][5], 4)
    res["pm_Phi"], res["pm_Gamma"] = Phi, Gam

    # LQR gains: run_simulation weights (run_simulation.py:54) and class defaults
    for tag, Qd, Rd in (("sim", [15.0, 15.0, 8.0], [0.1, 0.1]), ("def", [10.0, 10.0, 1.0], [0.1, 0.1]),
                        ("aniso", [30.0, 5.0, 2.0], [0.05, 0.4])):
        lqr = LQRController(Q_diag=Qd, R_diag=Rd, dt=0.02, v_max=2.0, omega_max=3.0)
        K = np.zeros((n, 2, 3)); P = np.zeros((n, 3, 3))
        for i in range(n):
            K[i] = lqr.compute_gain(v[i], th[i], force_recompute=True)
            P[i] = lqr.P
        res[f"K_{tag}"], res[f"P_{tag}"] = K, P
        res[f"Qd_{tag}"], res[f"Rd_{tag}"] = np.array(Qd), np.array(Rd)
    gen = Gen(A=2.0, a=0.5, dt=0.02)
    res["traj20"] = gen.generate(20.0)
    xs, us = gen.get_trajectory_segment(990, 21)
    res["seg990_x"], res["seg990_u"] = xs, us
    # LQR control law + plant (lqr_controller.py:149-215, differential_drive.py:138-172)
    lqr = LQRController(Q_diag=[15.0, 15.0, 8.0], R_diag=[0.1, 0.1], dt=0.02, v_max=2.0, omega_max=3.0)
    rob = Robot(v_max=2.0, omega_max=3.0)
    m = 64
    X = rng.uniform(-2, 2, (m, 3)); X[:, 2] = rng.uniform(-np.pi, np.pi, m)
    XR = X + rng.uniform(-0.3, 0.3, (m, 3)); XR[::7, 2] += 2 * np.pi
    UR = np.stack([rng.uniform(0.3, 1.5, m), rng.uniform(-1, 1, m)], 1)
    U = np.zeros((m, 2)); E = np.zeros((m, 3)); XN = np.zeros((m, 3))
    for i in range(m):
        U[i], E[i] = lqr.compute_control_at_operating_point(X[i].copy(), XR[i].copy(), UR[i].copy())
        XN[i] = rob.simulate_step(X[i].copy(), U[i] * 1.7, 0.02)
    res.update(law_x=X, law_xr=XR, law_ur=UR, law_u=U, law_e=E, plant_next=XN)
    np.savez_compressed(os.path.join(OUT, "ref_py.npz"), v=v, th=th, dt=0.02, **res)

    # risk supervisor (HC/controllers/risk_metrics.py) on seeded states, and the one hybrid run log the
    # reference ships (20260208_003249, run_hybrid_simulation): states / controls / controller tags
    from hybrid_controller.controllers.risk_metrics import RiskMetrics
    rm = RiskMetrics(d_safe=0.3, d_trigger=1.0, alpha=0.6, beta=0.4, threshold_low=0.2, threshold_medium=0.5)
    obs = [dict(x=1.0, y=0.5, radius=0.2), dict(x=-0.5, y=-1.0, radius=0.25), dict(x=1.5, y=-0.3, radius=0.15)]
    S = rng.uniform(-2.2, 2.2, (200, 3))
    pred = rng.uniform(-2.2, 2.2, (200, 6, 3)) * 0.5 + S[:, None, :]
    ra = np.zeros((200, 5)); rp = np.zeros(200)
    for i in range(200):
        a_ = rm.assess_risk(S[i], obs)
        ra[i] = [a_.distance_risk, a_.combined_risk, a_.min_obstacle_distance, a_.nearest_obstacle_id, float(a_.use_mpc)]
        rp[i] = rm.assess_risk(S[i], obs, pred[i]).combined_risk
    st, ct, _, _ = read_log("20260208_003249")
    tags = np.array([1 if r[3] == "MPC" else 0 for r in list(csv.reader(
        open(f"{REF}/logs/controls_20260208_003249.csv")))[1:]], dtype=np.int8)
    np.savez_compressed(os.path.join(OUT, "g_hyb.npz"), risk_states=S, risk_pred=pred, risk_out=ra, risk_pred_combined=rp,
                        risk_alpha=rm.alpha, risk_beta=rm.beta,
                        states=st[:, 1:4], controls=ct[:, 1:3], controller=tags,
                        source="logs/{states,controls}_20260208_003249.csv (hybrid_sim) + RiskMetrics on seeded states")
    print("wrote", sorted(os.listdir(OUT)))


if __name__ == "__main__":
    main()
