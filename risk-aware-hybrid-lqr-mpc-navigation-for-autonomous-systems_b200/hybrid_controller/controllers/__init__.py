"""controllers: LQRController (K2) and MPCController (K3); other controllers resolve from the
reference checkout when present (see package docstring)."""
from .. import _overlay
_overlay(__path__, "controllers")
from .lqr_controller import LQRController      # noqa: E402
from .mpc_controller import MPCController, Obstacle, MPCSolution  # noqa: E402
from .risk_metrics import RiskMetrics, RiskAssessment  # noqa: E402

__all__ = ["LQRController", "MPCController", "Obstacle", "MPCSolution", "RiskMetrics", "RiskAssessment"]
