from .reference_generator import ReferenceTrajectoryGenerator

__all__ = ["ReferenceTrajectoryGenerator"]
