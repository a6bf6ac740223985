#!/usr/bin/env python3
"""Run the reference's UNMODIFIED run_simulation.py on top of the B200 drop-in package.

    python tools/run_reference_sim.py --mode lqr --no-plot        (needs a GPU and a reference checkout:
                                                                   $REF_ROOT, /root/reference or baseline/_ref)
The drop-in `hybrid_controller` is imported first, so the script's own `sys.path.insert` of the reference
package cannot shadow it; out-of-scope modules (logger, risk metrics, ...) resolve from the reference
through the package overlay.  matplotlib is absent from this image: a stub keeps the import of the
reference's visualization module working (use --no-plot)."""
import os
import runpy
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "risk-aware-hybrid-lqr-mpc-navigation-for-autonomous-systems_b200"))

class _Stub:
    """Permissive stand-in for any matplotlib object (graphics are unavailable on this box; --no-plot)."""
    def __init__(self, *a, **k): pass
    def __call__(self, *a, **k): return _Stub()
    def __getattr__(self, name): return _Stub()
    def __iter__(self): return iter(())


try:
    import matplotlib  # noqa: F401
    import matplotlib.pyplot  # noqa: F401
except Exception:
    for name in ("matplotlib", "matplotlib.pyplot", "matplotlib.patches", "matplotlib.collections"):
        m = types.ModuleType(name)
        m.__getattr__ = lambda attr: _Stub()        # `from matplotlib.patches import Circle` -> a dummy callable
        m.__path__ = []
        sys.modules[name] = m
    for sub in ("pyplot", "patches", "collections"):
        setattr(sys.modules["matplotlib"], sub, sys.modules["matplotlib." + sub])
    sys.modules["matplotlib"].use = lambda *a, **k: None

import hybrid_controller  # noqa: E402,F401  (the drop-in; must win over the reference's package)
import hybrid_controller.controllers  # noqa: E402,F401
import hybrid_controller.models  # noqa: E402,F401
import hybrid_controller.trajectory  # noqa: E402,F401

REF = hybrid_controller.reference_root()
if REF is None:
    raise SystemExit("no reference checkout ($REF_ROOT, /root/reference or baseline/_ref): run tools/stage_reference.py")
print(f"[run_reference_sim] reference: {REF}   drop-in: {os.path.dirname(hybrid_controller.__file__)}", flush=True)
sys.argv = [os.path.join(REF, "run_simulation.py")] + sys.argv[1:]
runpy.run_path(sys.argv[0], run_name="__main__")
