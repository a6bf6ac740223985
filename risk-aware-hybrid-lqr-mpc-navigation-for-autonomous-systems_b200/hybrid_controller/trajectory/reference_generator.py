"""Host mirror of the figure-8 reference generator (the input generator of every
benchmark configuration; it must exist on the GPU box, where /root/reference does not).

Same public surface and numerics as /root/reference/.../trajectory/reference_generator.py (:42-350):
px = A sin(at), py = A sin(at) cos(at); heading = atan2 of the analytic velocity; omega is the
FORWARD DIFFERENCE of the heading over dt with wrap (Q16); lookups clamp at the last row.
Scalar evaluation order is kept so generate() reproduces the reference's table bit for bit
(checked against tests/golden/ref_py.npz).
"""
from typing import Optional, Tuple

import numpy as np


class ReferenceTrajectoryGenerator:
    def __init__(self, A: float = 2.0, a: float = 0.5, dt: float = 0.02):
        self.A, self.a, self.dt = A, a, dt
        self._trajectory: Optional[np.ndarray] = None
        self._duration: float = 0.0

    # -- analytic curve --------------------------------------------------------------------------
    def position(self, t: float) -> Tuple[float, float]:
        s = np.sin(self.a * t)
        return self.A * s, self.A * s * np.cos(self.a * t)

    def velocity(self, t: float) -> Tuple[float, float]:
        c, s = np.cos(self.a * t), np.sin(self.a * t)
        return self.a * self.A * c, self.a * self.A * (c**2 - s**2)

    def heading(self, t: float) -> float:
        dpx, dpy = self.velocity(t)
        return np.arctan2(dpy, dpx)

    def linear_velocity(self, t: float) -> float:
        dpx, dpy = self.velocity(t)
        return np.sqrt(dpx**2 + dpy**2)

    def angular_velocity(self, t: float) -> float:
        d = self.heading(t + self.dt) - self.heading(t)
        while d > np.pi:
            d -= 2 * np.pi
        while d < -np.pi:
            d += 2 * np.pi
        return d / self.dt

    def get_reference_at_time(self, t: float) -> Tuple[np.ndarray, np.ndarray]:
        px, py = self.position(t)
        return (np.array([px, py, self.heading(t)]),
                np.array([self.linear_velocity(t), self.angular_velocity(t)]))

    # -- sampled table ---------------------------------------------------------------------------
    def generate(self, duration: float) -> np.ndarray:
        """(n,6) rows [t, px, py, theta, v, omega], t = arange(0, duration, dt)."""
        ts = np.arange(0, duration, self.dt)
        table = np.zeros((len(ts), 6))
        for k, tk in enumerate(ts):
            x_ref, u_ref = self.get_reference_at_time(tk)
            table[k, 0] = tk
            table[k, 1:4] = x_ref
            table[k, 4:6] = u_ref
        self._trajectory, self._duration = table, duration
        return table

    def generate_figure_8(self, t_range: float, dt: float, A: float = None, a: float = None) -> np.ndarray:
        """Vectorised (n,5) variant [x, y, theta, v, omega] with np.diff omega (last omega = 0)."""
        A = self.A if A is None else A
        a = self.a if a is None else a
        t = np.arange(0, t_range, dt)
        s, c = np.sin(a * t), np.cos(a * t)
        dxr, dyr = a * A * c, a * A * (c**2 - s**2)
        th = np.arctan2(dyr, dxr)
        w = np.diff(th, append=th[-1]) / dt
        w = np.where(w > np.pi / dt, w - 2 * np.pi / dt, w)
        w = np.where(w < -np.pi / dt, w + 2 * np.pi / dt, w)
        return np.stack([A * s, A * s * c, th, np.sqrt(dxr**2 + dyr**2), w], axis=1)

    def _need_table(self) -> np.ndarray:
        if self._trajectory is None:
            raise ValueError("Trajectory not generated. Call generate() first.")
        return self._trajectory

    def get_reference_at_index(self, k: int) -> Tuple[np.ndarray, np.ndarray]:
        table = self._need_table()
        row = table[min(k, len(table) - 1)]
        return row[1:4].copy(), row[4:6].copy()

    def get_trajectory_segment(self, start_idx: int, horizon: int) -> Tuple[np.ndarray, np.ndarray]:
        table = self._need_table()
        idx = np.minimum(start_idx + np.arange(horizon), len(table) - 1)
        return table[idx, 1:4].copy(), table[idx, 4:6].copy()

    @property
    def num_points(self) -> int:
        return 0 if self._trajectory is None else len(self._trajectory)

    def get_trajectory_bounds(self) -> Tuple[Tuple[float, float], Tuple[float, float]]:
        if self._trajectory is None:
            return ((-self.A, self.A), (-self.A / 2, self.A / 2))
        t = self._trajectory
        return ((t[:, 1].min(), t[:, 1].max()), (t[:, 2].min(), t[:, 2].max()))
