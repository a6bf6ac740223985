"""Drop-in ``LQRController`` backed by the batched fp64 DARE kernel (K2).

Mirrors /root/reference/src/hybrid_controller/hybrid_controller/controllers/lqr_controller.py
(:31-283): constructor signature and defaults, public attributes (K, P, Q, R, dt, v_max,
omega_max, linearizer, _last_v_r, _last_theta_r), the gain cache, the |v|<1e-6 guard, the
fallback gain with its printed warning, the wrap/clip helpers.  scipy's solve_discrete_are is
replaced by the structured-doubling kernel in libb2c.so (linearization fused in); there is no
CPU fallback.  ``*_batch`` methods are additions.
"""
from typing import Optional, Tuple

import numpy as np

from .. import _native as nat
from ..models.linearization import Linearizer


class LQRController:
    def __init__(self, Q_diag: list = None, R_diag: list = None,
                 dt: float = 0.02, v_max: float = 1.0, omega_max: float = 1.5):
        # defaults: lqr_controller.py:69-73
        if Q_diag is None:
            Q_diag = [10.0, 10.0, 1.0]
        if R_diag is None:
            R_diag = [0.1, 0.1]
        self.Q = np.diag(Q_diag)
        self.R = np.diag(R_diag)
        self.dt = dt
        self.v_max = v_max
        self.omega_max = omega_max
        self.linearizer = Linearizer(dt=dt)
        self.K: Optional[np.ndarray] = None
        self.P: Optional[np.ndarray] = None
        self._last_v_r: float = 0.0
        self._last_theta_r: float = 0.0

    @property
    def _ctx(self):
        return nat.default_context(-1)

    def _weights(self):
        Q, R = np.asarray(self.Q, float), np.asarray(self.R, float)
        if np.any(Q != np.diag(np.diag(Q))) or np.any(R != np.diag(np.diag(R))):
            raise NotImplementedError("the batched DARE kernel supports diagonal Q and R "
                                      "(the only form the constructor/set_weights can produce)")
        return np.diag(Q), np.diag(R)

    # -- reference API ---------------------------------------------------------------------------
    def compute_gain(self, v_r: float, theta_r: float, force_recompute: bool = False) -> np.ndarray:
        """lqr_controller.py:92-147."""
        if not force_recompute and self.K is not None:
            if abs(v_r - self._last_v_r) < 1e-6 and abs(theta_r - self._last_theta_r) < 1e-6:
                return self.K
        Qd, Rd = self._weights()
        K, P, _, st = nat.dare_host(self._ctx, [float(v_r)], [float(theta_r)], self.dt, Qd, Rd,
                                    nat.LIN_EULER, nat.GUARD_LQR, want_P=True)
        if st[0] == nat.ST_OK:
            self.P = P[0]
            self.K = K[0]
        else:
            # kernel already wrote the reference's fallback gain [[1,0,0],[0,0,1]] (:134-141)
            self.K = K[0]
            reason = {nat.ST_MAXITER: "doubling iteration did not converge (no stabilizing solution)",
                      nat.ST_NONFINITE: "non-finite Riccati iterate"}.get(int(st[0]), f"status {st[0]}")
            print(f"Warning: DARE solver failed, using fallback gain. Error: {reason}")
        self._last_v_r = v_r
        self._last_theta_r = theta_r
        return self.K

    def compute_control(self, x: np.ndarray, x_ref: np.ndarray,
                        u_ref: np.ndarray, K: np.ndarray = None) -> np.ndarray:
        """lqr_controller.py:149-189.  2x3 matvec + wrap + clip for ONE robot: host arithmetic
        (bit-identical to the reference); the batched law is compute_control_batch (kernel)."""
        if K is None:
            if self.K is None:
                self.compute_gain(u_ref[0], x_ref[2])
            K = self.K
        error = x - x_ref
        error[2] = self._normalize_angle(error[2])
        u_feedback = -K @ error
        u = u_ref + u_feedback
        return self._clip_control(u)

    def compute_control_at_operating_point(self, x: np.ndarray, x_ref: np.ndarray,
                                           u_ref: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
        """lqr_controller.py:191-215."""
        K = self.compute_gain(u_ref[0], x_ref[2])
        error = x - x_ref
        error[2] = self._normalize_angle(error[2])
        u = self.compute_control(x, x_ref, u_ref, K)
        return u, error

    def get_lqr_gain(self, v_r: float, theta_r: float, dt: float = None) -> np.ndarray:
        """lqr_controller.py:217-242: uncached, no low-speed guard, raises if the DARE has no
        solution (scipy raises LinAlgError there), does not touch self.K / self.P."""
        use_dt = dt if (dt is not None and abs(dt - self.dt) > 1e-9) else self.dt
        Qd, Rd = self._weights()
        K, _, _, st = nat.dare_host(self._ctx, [float(v_r)], [float(theta_r)], use_dt, Qd, Rd,
                                    nat.LIN_EULER, nat.GUARD_NONE, want_P=False)
        if st[0] != nat.ST_OK:
            raise np.linalg.LinAlgError("Failed to find a finite solution.")
        return K[0]

    def _normalize_angle(self, angle: float) -> float:
        """lqr_controller.py:244-250."""
        while angle > np.pi:
            angle -= 2 * np.pi
        while angle < -np.pi:
            angle += 2 * np.pi
        return angle

    def _clip_control(self, u: np.ndarray) -> np.ndarray:
        """lqr_controller.py:252-257."""
        return np.array([
            np.clip(u[0], -self.v_max, self.v_max),
            np.clip(u[1], -self.omega_max, self.omega_max)
        ])

    def get_cost_matrices(self) -> Tuple[np.ndarray, np.ndarray]:
        return self.Q.copy(), self.R.copy()

    def set_weights(self, Q_diag: list = None, R_diag: list = None) -> None:
        """lqr_controller.py:263-278."""
        if Q_diag is not None:
            self.Q = np.diag(Q_diag)
        if R_diag is not None:
            self.R = np.diag(R_diag)
        self.K = None
        self.P = None

    @property
    def gain_computed(self) -> bool:
        return self.K is not None

    # -- batch additions ---------------------------------------------------------------------------
    def compute_gain_batch(self, v_r, theta_r, want_P: bool = False, guard: bool = True):
        """(B,), (B,) -> K (B,2,3) [, P (B,3,3)], status (B,) int32.  No cache.
        Instances whose DARE fails carry the reference's fallback gain and status != 0."""
        Qd, Rd = self._weights()
        K, P, it, st = nat.dare_host(self._ctx, v_r, theta_r, self.dt, Qd, Rd, nat.LIN_EULER,
                                     nat.GUARD_LQR if guard else nat.GUARD_NONE, want_P=want_P)
        return (K, P, st) if want_P else (K, st)

    def compute_control_batch(self, x, x_ref, u_ref):
        """Batched compute_control_at_operating_point: x (B,3), x_ref (B,3), u_ref (B,2) ->
        u (B,2), wrapped error (B,3).  DARE (K2) + law on the device, host numpy in and out."""
        x = np.ascontiguousarray(x, float).reshape(-1, 3)
        B = x.shape[0]
        x_ref = np.ascontiguousarray(x_ref, float).reshape(B, 3)
        u_ref = np.ascontiguousarray(u_ref, float).reshape(B, 2)
        K, _ = self.compute_gain_batch(u_ref[:, 0], x_ref[:, 2])          # compute_gain(u_ref[0], x_ref[2]), :206
        u, e = np.empty((B, 2)), np.empty((B, 3))
        self._ctx.call("b2c_lqr_control_host", B, float(self.v_max), float(self.omega_max), nat._ptr(x), nat._ptr(x_ref),
                       nat._ptr(u_ref), nat._ptr(np.ascontiguousarray(K)), nat._ptr(u), nat._ptr(e))
        return u, e
