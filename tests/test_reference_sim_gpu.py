"""SURVEY 8(b) acceptance test: the reference's UNMODIFIED run_simulation.py (main :641-671, run_lqr_simulation :34-136,
run_mpc_simulation :139-335) runs on top of the drop-in package -- import overlay, SimulationLogger hand-off
(:83-84, :267), mpc.N / mpc.d_safe reads (:247, :296, :312) -- and the CSVs the reference's own SimulationLogger writes
reproduce the shipped run logs (tests/golden/g_lqr.npz, g_mpc.npz) within the bands of conftest.py.

Needs a reference checkout: $REF_ROOT, /root/reference, or the staged copy baseline/_ref (tools/stage_reference.py,
run by __graft_entry__.build(); git-ignored, travels to the GPU box)."""
import csv
import glob
import os
import re
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT, assert_g_mpc_closed_loop

pytestmark = pytest.mark.gpu


def _run(mode, tmp_path):
    import hybrid_controller
    ref = hybrid_controller.reference_root()
    if ref is None:
        pytest.skip("no reference checkout ($REF_ROOT, /root/reference, baseline/_ref): run tools/stage_reference.py")
    cmd = [sys.executable, os.path.join(ROOT, "tools", "run_reference_sim.py"), "--mode", mode, "--no-plot"]
    r = subprocess.run(cmd, cwd=str(tmp_path), capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "drop-in:" in r.stdout and "risk-aware-hybrid" in r.stdout           # the drop-in, not the reference's classes

    def load(kind):
        files = glob.glob(os.path.join(str(tmp_path), "logs", f"{kind}_*.csv"))
        assert len(files) == 1, files
        with open(files[0]) as f:
            return list(csv.DictReader(f))
    return r.stdout, load("states"), load("controls")


def test_reference_run_simulation_lqr_mode_unmodified(golden, tmp_path):
    out, st, ct = _run("lqr", tmp_path)
    g = golden["g_lqr"]
    x = np.array([[float(r[k]) for k in ("px", "py", "theta")] for r in st])
    u = np.array([[float(r[k]) for k in ("v", "omega")] for r in ct])
    assert x.shape == (999, 3) and u.shape == (999, 2) and all(r["controller"] == "LQR" for r in ct)
    assert np.abs(x - g["states"][:999]).max() <= 1e-9 and np.abs(u - g["controls"][:999]).max() <= 1e-9
    e = np.array([[float(r[k]) for k in ("error_px", "error_py", "error_theta")] for r in st])
    # summary of logs/simulation_20260207_234542.log:2 (3-vector error norm over the 999 logged ticks)
    en = np.linalg.norm(e, axis=1)
    assert abs(en.mean() - 0.007625199505240531) < 1e-9 and abs(en.max() - 0.012787467899794236) < 1e-9
    m = re.search(r"Mean tracking error: ([0-9.]+) m", out)
    assert m and abs(float(m.group(1)) - np.linalg.norm(g["errors"][:999, :2], axis=1).mean()) < 1e-4


def test_reference_run_simulation_mpc_mode_unmodified(golden, tmp_path):
    out, st, ct = _run("mpc", tmp_path)
    g = golden["g_mpc"]
    x = np.array([[float(r[k]) for k in ("px", "py", "theta")] for r in st])
    u = np.array([[float(r[k]) for k in ("v", "omega")] for r in ct])
    assert x.shape == (999, 3) and u.shape == (999, 2) and all(r["controller"] == "MPC" for r in ct)
    assert_g_mpc_closed_loop(x, u, g["states"], g["controls"])
    ms = np.array([float(r["solve_time_ms"]) for r in ct])
    assert np.median(ms) < 5.0                                                 # shipped log: 82.6 ms mean
    assert re.search(r"Collision events: \d+", out) and "Mean MPC solve time" in out


def test_reference_run_simulation_hybrid_mode_unmodified(tmp_path):
    """`--mode hybrid` (run_hybrid_simulation :413-638): the reference's own loop drives the drop-in LQRController,
    MPCController (block_size 1, every tick) and RiskMetrics mirror, one scalar call per tick.  The reference ships no
    log of this mode with its final parameters, so the run is compared with the DEVICE rollout engine executing the same
    loop for the same robot (b2c_rollout_host, hybrid mode): identical controller choice at every tick, states to 1e-9."""
    import re as _re
    from hybrid_controller.rollout import simulate_batch
    from hybrid_controller.controllers import LQRController, MPCController, RiskMetrics
    from hybrid_controller import workloads as wl
    out, st, ct = _run("hybrid", tmp_path)
    x = np.array([[float(r[k]) for k in ("px", "py", "theta")] for r in st])
    u = np.array([[float(r[k]) for k in ("v", "omega")] for r in ct])
    tags = np.array([1 if r["controller"] == "MPC" else 0 for r in ct])
    assert x.shape == (999, 3) and 0 < tags.sum() < 999
    m = _re.search(r"LQR steps: (\d+).*\n\s*MPC steps: (\d+).*\n\s*Controller switches: (\d+)", out)
    assert m and int(m.group(1)) + int(m.group(2)) == 999
    table = wl.figure8_table()
    ref = simulate_batch("hybrid", table, [0], table[0:1, 1:4], wl.DEFAULT_OBSTACLES, 999, lqr=LQRController(**wl.SIM_LQR),
                         mpc=MPCController(horizon=6, block_size=1, **wl.SIM_MPC),
                         risk=RiskMetrics(d_safe=0.3, d_trigger=1.0, alpha=0.6, beta=0.4, threshold_low=0.2,
                                          threshold_medium=0.5), log=True)
    assert np.array_equal(ref["controller"][0], tags)
    assert int(m.group(1)) == ref["lqr_steps"][0] and int(m.group(2)) == ref["mpc_solves"][0]
    assert int(m.group(3)) == ref["switches"][0]
    assert np.abs(ref["states"][0][:999] - x).max() <= 1e-9 and np.abs(ref["controls"][0] - u).max() <= 1e-9
