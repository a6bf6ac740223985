"""B200-native drop-in for the hot path of the reference's ``hybrid_controller`` package.

Built here (GPU-backed through libb2c.so): models.linearization.Linearizer,
controllers.lqr_controller.LQRController, controllers.mpc_controller.{MPCController, Obstacle,
MPCSolution}; host-side mirrors of the input generator and the plant
(trajectory.reference_generator, models.differential_drive).

Everything else of the reference package (logging, utils, risk_metrics, yaw_stabilizer, nodes) is
out of scope and is NOT re-implemented: when the reference checkout is present
(``$REF_ROOT``, /root/reference, or the staged copy ``<repo>/baseline/_ref`` that travels to the GPU box) its
directories are appended to the sub-package search paths, so
the reference's own ``run_simulation.py`` runs unmodified on top of this package and picks those
modules up from the reference itself.
"""
import os as _os

__version__ = "0.1.0"


def reference_root():
    """Directory of the reference checkout (holds run_simulation.py), or None."""
    here = _os.path.dirname(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
    for root in (_os.environ.get("REF_ROOT"), "/root/reference", _os.path.join(here, "baseline", "_ref")):
        if root and _os.path.isfile(_os.path.join(root, "run_simulation.py")):
            return root
    return None


def _reference_pkg_dir():
    for root in (reference_root(),):
        if root:
            d = _os.path.join(root, "src", "hybrid_controller", "hybrid_controller")
            if _os.path.isdir(d):
                return d
    return None


def _overlay(path_list, sub=""):
    ref = _reference_pkg_dir()
    if ref:
        d = _os.path.join(ref, sub) if sub else ref
        if _os.path.isdir(d) and d not in path_list:
            path_list.append(d)


_overlay(__path__)
