import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "risk-aware-hybrid-lqr-mpc-navigation-for-autonomous-systems_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def golden():
    import numpy as np
    return {k: np.load(os.path.join(GOLDEN, k + ".npz")) for k in ("g_lqr", "g_mpc", "ref_py")}


# run_simulation.py parameter sets
MPC_PRM6 = dict(N=6, Qd=[15, 15, 50], Rd=[.1, .1], Pd=[30, 30, 40], d_safe=.3, rho=5000., v_max=2.,
                w_max=3., dt=.02, block_size=2)
MPC_PRM20 = dict(MPC_PRM6, N=20, block_size=1)
DEFAULT_OBS = [(1.0, .5, .2), (-.5, -1.0, .25), (1.5, -.3, .15)]
