"""models: Linearizer (K1, GPU) and the differential-drive plant (host mirror)."""
from .. import _overlay
_overlay(__path__, "models")
from .linearization import Linearizer          # noqa: E402
from .differential_drive import DifferentialDriveRobot  # noqa: E402

__all__ = ["Linearizer", "DifferentialDriveRobot"]
