"""Host mirror of the unicycle plant (input to the closed-loop configs).

Same public surface as /root/reference/.../models/differential_drive.py (:33-279):
RobotState, ControlInput, DifferentialDriveRobot.{continuous_dynamics, simulate_step,
simulate_trajectory, clip_control, normalize_angle, compute_tracking_error, get_wheel_velocities,
from_wheel_velocities}.  Single-robot scalar arithmetic in the same operation order, so closed
loops replay the shipped run logs; the batched plant step lives in the rollout kernel.
"""
from dataclasses import dataclass
from typing import Tuple

import numpy as np

_TWO_PI = 2 * np.pi


@dataclass
class RobotState:
    px: float
    py: float
    theta: float

    def to_array(self) -> np.ndarray:
        return np.array([self.px, self.py, self.theta])

    @classmethod
    def from_array(cls, arr: np.ndarray) -> "RobotState":
        return cls(px=arr[0], py=arr[1], theta=arr[2])

    def __repr__(self) -> str:
        return f"RobotState(px={self.px:.3f}, py={self.py:.3f}, theta={self.theta:.3f})"


@dataclass
class ControlInput:
    v: float
    omega: float

    def to_array(self) -> np.ndarray:
        return np.array([self.v, self.omega])

    @classmethod
    def from_array(cls, arr: np.ndarray) -> "ControlInput":
        return cls(v=arr[0], omega=arr[1])

    def __repr__(self) -> str:
        return f"ControlInput(v={self.v:.3f}, omega={self.omega:.3f})"


class DifferentialDriveRobot:
    STATE_DIM = 3
    CONTROL_DIM = 2

    def __init__(self, v_max: float = 1.0, omega_max: float = 1.5, wheel_base: float = 0.3):
        self.v_max, self.omega_max, self.wheel_base = v_max, omega_max, wheel_base

    def continuous_dynamics(self, state: np.ndarray, control: np.ndarray) -> np.ndarray:
        """f(x,u) = [v cos(theta), v sin(theta), omega]  (differential_drive.py:111-136)."""
        theta = state[2]
        v, omega = control
        return np.array([v * np.cos(theta), v * np.sin(theta), omega])

    def simulate_step(self, state: np.ndarray, control: np.ndarray, dt: float,
                      method: str = 'euler') -> np.ndarray:
        """clip -> integrate -> wrap theta  (differential_drive.py:138-172)."""
        u = self.clip_control(control)
        f = self.continuous_dynamics
        if method == 'euler':
            nxt = state + dt * f(state, u)
        elif method == 'rk4':
            k1 = f(state, u)
            k2 = f(state + 0.5 * dt * k1, u)
            k3 = f(state + 0.5 * dt * k2, u)
            k4 = f(state + dt * k3, u)
            nxt = state + (dt / 6.0) * (k1 + 2 * k2 + 2 * k3 + k4)
        else:
            raise ValueError(f"Unknown integration method: {method}")
        nxt[2] = self.normalize_angle(nxt[2])
        return nxt

    def simulate_trajectory(self, x0: np.ndarray, controls: np.ndarray, dt: float,
                            method: str = 'euler') -> np.ndarray:
        out = np.zeros((len(controls) + 1, self.STATE_DIM))
        out[0] = np.array(x0, dtype=float)
        for k, u in enumerate(controls):
            out[k + 1] = self.simulate_step(out[k], u, dt, method)
        return out

    def clip_control(self, control: np.ndarray) -> np.ndarray:
        return np.array([np.clip(control[0], -self.v_max, self.v_max),
                         np.clip(control[1], -self.omega_max, self.omega_max)])

    @staticmethod
    def normalize_angle(angle: float) -> float:
        """closed interval [-pi, pi] by +-2pi steps (differential_drive.py:215-230)."""
        while angle > np.pi:
            angle -= _TWO_PI
        while angle < -np.pi:
            angle += _TWO_PI
        return angle

    def compute_tracking_error(self, state: np.ndarray, state_ref: np.ndarray) -> np.ndarray:
        err = state - state_ref
        err[2] = self.normalize_angle(err[2])
        return err

    def get_wheel_velocities(self, v: float, omega: float) -> Tuple[float, float]:
        half = self.wheel_base / 2
        return v - half * omega, v + half * omega

    def from_wheel_velocities(self, v_left: float, v_right: float) -> Tuple[float, float]:
        return (v_right + v_left) / 2, (v_right - v_left) / self.wheel_base
