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
