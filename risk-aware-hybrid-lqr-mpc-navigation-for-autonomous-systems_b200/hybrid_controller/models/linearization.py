"""Drop-in ``Linearizer`` backed by the batched sm_100a linearization kernel (K1).

Mirrors /root/reference/src/hybrid_controller/hybrid_controller/models/linearization.py
(class Linearizer, :32-321): same method names, argument meaning, return shapes/dtypes and the
``ValueError`` for an unknown ``method``.  Every model evaluation runs on the GPU through
libb2c.so (no CPU fallback); the ``*_batch`` methods are additions for the data-parallel path.
"""
from typing import Tuple

import numpy as np

from .. import _native as nat


class Linearizer:
    STATE_DIM = 3    # [px, py, theta]      (linearization.py:50)
    CONTROL_DIM = 2  # [v, omega]           (linearization.py:51)

    def __init__(self, dt: float = 0.02, device: int = -1):
        self.dt = dt
        self._device = device

    # -- engine access -------------------------------------------------------------------------
    @property
    def _ctx(self):
        return nat.default_context(self._device)

    def _one(self, v_r, theta_r, method):
        A, B = nat.linearize_host(self._ctx, [float(v_r)], [float(theta_r)], self.dt, method)
        return A[0], B[0]

    # -- reference API ---------------------------------------------------------------------------
    def get_jacobians(self, v_r: float, theta_r: float) -> Tuple[np.ndarray, np.ndarray]:
        """Continuous-time Jacobians (linearization.py:62-96)."""
        return self._one(v_r, theta_r, nat.LIN_JACOBIAN)

    def discretize_euler(self, A: np.ndarray, B: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
        """A_d = I + A T_s, B_d = B T_s (linearization.py:98-118); 15 flops on given matrices --
        host-side glue, the operating-point evaluation is what the kernel batches."""
        return np.eye(self.STATE_DIM) + np.asarray(A) * self.dt, np.asarray(B) * self.dt

    def discretize_exact(self, A: np.ndarray, B: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
        """Exact ZOH (linearization.py:120-158).  For this model A is nilpotent (A@A == 0), so
        expm terminates: A_d = I + A T_s, B_d = (I T_s + A T_s^2/2) B -- the closed form K1 uses."""
        A = np.asarray(A, dtype=float)
        B = np.asarray(B, dtype=float)
        dt = self.dt
        if np.any(A @ A != 0.0):
            # not a Jacobian this model can produce: generic augmented expm, as the reference does
            from scipy.linalg import expm
            n, m = A.shape[0], B.shape[1]
            aug = np.zeros((n + m, n + m))
            aug[:n, :n] = A * dt
            aug[:n, n:] = B * dt
            return expm(A * dt), expm(aug)[:n, n:]
        return np.eye(self.STATE_DIM) + A * dt, (np.eye(self.STATE_DIM) * dt + A * (dt * dt / 2)) @ B

    def get_discrete_model(self, v_r: float, theta_r: float,
                           method: str = 'euler') -> Tuple[np.ndarray, np.ndarray]:
        """linearization.py:160-188."""
        if method == 'euler':
            return self._one(v_r, theta_r, nat.LIN_EULER)
        elif method == 'exact':
            return self._one(v_r, theta_r, nat.LIN_EXACT)
        else:
            raise ValueError(f"Unknown discretization method: {method}")

    def get_discrete_model_explicit(self, v_r: float, theta_r: float) -> Tuple[np.ndarray, np.ndarray]:
        """linearization.py:190-225 (identical to method='euler')."""
        return self._one(v_r, theta_r, nat.LIN_EULER)

    def predict_trajectory(self, x0: np.ndarray, controls: np.ndarray,
                           v_refs: np.ndarray, theta_refs: np.ndarray) -> np.ndarray:
        """LTV open-loop rollout (linearization.py:227-255): one batched K1 call for all stages."""
        controls = np.asarray(controls, dtype=float)
        N = len(controls)
        trajectory = np.zeros((N + 1, self.STATE_DIM))
        trajectory[0] = np.asarray(x0, dtype=float)
        if N == 0:
            return trajectory
        Ad, Bd = self.get_discrete_model_explicit_batch(np.asarray(v_refs, float)[:N],
                                                        np.asarray(theta_refs, float)[:N])
        for k in range(N):
            trajectory[k + 1] = Ad[k] @ trajectory[k] + Bd[k] @ controls[k]
        return trajectory

    def predict_horizon(self, x0: np.ndarray, u_seq: np.ndarray,
                        v_r: float, theta_r: float) -> np.ndarray:
        """LTI open-loop rollout (linearization.py:257-280)."""
        A_d, B_d = self.get_discrete_model_explicit(v_r, theta_r)
        u_seq = np.asarray(u_seq, dtype=float)
        N = len(u_seq)
        trajectory = np.zeros((N + 1, self.STATE_DIM))
        trajectory[0] = np.asarray(x0, dtype=float)
        for k in range(N):
            trajectory[k + 1] = A_d @ trajectory[k] + B_d @ u_seq[k]
        return trajectory

    @staticmethod
    def build_prediction_matrices(A_d: np.ndarray, B_d: np.ndarray, N: int) -> Tuple[np.ndarray, np.ndarray]:
        """Condensed X = Phi x0 + Gamma U (linearization.py:282-321).  Not per-instance work and
        unused by every caller in the reference: plain host numpy, powers built incrementally."""
        n, m = A_d.shape[0], B_d.shape[1]
        powers = [np.eye(n)]
        for _ in range(N):
            powers.append(powers[-1] @ A_d)
        Phi = np.vstack(powers[1:]) if N else np.zeros((0, n))
        Gamma = np.zeros((n * N, m * N))
        for i in range(N):
            for j in range(i + 1):
                Gamma[i * n:(i + 1) * n, j * m:(j + 1) * m] = powers[i - j] @ B_d
        return Phi, Gamma

    # -- batch additions ---------------------------------------------------------------------------
    def get_discrete_model_batch(self, v_r, theta_r, method: str = 'euler'):
        """(B,) , (B,) -> A_d (B,3,3), B_d (B,3,2); method in {'euler','exact','jacobian'}."""
        ids = {'euler': nat.LIN_EULER, 'exact': nat.LIN_EXACT, 'jacobian': nat.LIN_JACOBIAN}
        if method not in ids:
            raise ValueError(f"Unknown discretization method: {method}")
        return nat.linearize_host(self._ctx, v_r, theta_r, self.dt, ids[method])

    def get_discrete_model_explicit_batch(self, v_r, theta_r):
        return nat.linearize_host(self._ctx, v_r, theta_r, self.dt, nat.LIN_EULER)

    def get_jacobians_batch(self, v_r, theta_r):
        return nat.linearize_host(self._ctx, v_r, theta_r, self.dt, nat.LIN_JACOBIAN)
