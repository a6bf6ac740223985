#!/usr/bin/env python3
"""Run the reference's UNMODIFIED run_simulation.py on top of the B200 drop-in package.

    python tools/run_reference_sim.py --mode lqr --no-plot        (needs a GPU and /root/reference)
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
REF = os.environ.get("REF_ROOT", "/root/reference")

for name in ("matplotlib", "matplotlib.pyplot", "matplotlib.patches", "matplotlib.collections"):
    try:
        __import__(name)
    except Exception:
        m = types.ModuleType(name)
        m.__getattr__ = lambda attr: (_ for _ in ()).throw(AttributeError(attr))
        sys.modules[name] = m
if not hasattr(sys.modules["matplotlib"], "use"):
    sys.modules["matplotlib"].use = lambda *a, **k: None

import hybrid_controller  # noqa: E402,F401  (the drop-in; must win over the reference's package)
import hybrid_controller.controllers  # noqa: E402,F401
import hybrid_controller.models  # noqa: E402,F401
import hybrid_controller.trajectory  # noqa: E402,F401

sys.argv = [os.path.join(REF, "run_simulation.py")] + sys.argv[1:]
runpy.run_path(sys.argv[0], run_name="__main__")
