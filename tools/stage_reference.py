#!/usr/bin/env python3
"""Stage the parts of the reference checkout that the acceptance run of SURVEY 8(b) needs into the git-ignored
``baseline/_ref/`` (it travels to the GPU box with the snapshot; /root/reference does not exist there):

    run_simulation.py                               the UNMODIFIED driver (configs 1, 2 of BASELINE.json)
    src/hybrid_controller/hybrid_controller/**.py   the reference package (logger, visualisation, yaw stabiliser, ...
                                                    are taken from it through the drop-in package's path overlay)

Nothing under baseline/_ref is product source or committed; the shipped run logs the run is diffed against are the
committed fixtures tests/golden/g_lqr.npz / g_mpc.npz (tools/make_golden.py)."""
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def stage(src="/root/reference", dst=None):
    dst = dst or os.path.join(ROOT, "baseline", "_ref")
    if not os.path.isdir(src):
        return None
    os.makedirs(dst, exist_ok=True)
    shutil.copy2(os.path.join(src, "run_simulation.py"), os.path.join(dst, "run_simulation.py"))
    pkg_src = os.path.join(src, "src", "hybrid_controller")
    pkg_dst = os.path.join(dst, "src", "hybrid_controller")
    if os.path.isdir(pkg_dst):
        shutil.rmtree(pkg_dst)
    shutil.copytree(pkg_src, pkg_dst, ignore=shutil.ignore_patterns("__pycache__", "*.pyc"))
    return dst


if __name__ == "__main__":
    out = stage(*(sys.argv[1:2]))
    print(out or "no reference checkout found")
