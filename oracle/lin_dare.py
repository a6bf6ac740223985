"""Oracle: linearization + DARE/LQR gain (numpy / scipy, fp64).  TEST INFRASTRUCTURE.

Citations are to /root/reference/src/hybrid_controller/hybrid_controller/ (HC/).
The DARE arithmetic itself is scipy.linalg.solve_discrete_are -- the very routine
the reference calls (HC/controllers/lqr_controller.py:126,237) -- so this oracle is
the reference's own numerics, not an approximation of them.
"""
import numpy as np
from scipy.linalg import expm, solve_discrete_are


def jacobians(v, th):
    """HC/models/linearization.py:62-96."""
    A = np.array([[0, 0, -v * np.sin(th)], [0, 0, v * np.cos(th)], [0, 0, 0]], dtype=float)
    B = np.array([[np.cos(th), 0], [np.sin(th), 0], [0, 1]], dtype=float)
    return A, B


def discrete_explicit(v, th, dt):
    """HC/models/linearization.py:190-225 (first-order 'ZOH')."""
    s, c = np.sin(th), np.cos(th)
    Ad = np.array([[1, 0, -v * s * dt], [0, 1, v * c * dt], [0, 0, 1]], dtype=float)
    Bd = np.array([[c * dt, 0], [s * dt, 0], [0, dt]], dtype=float)
    return Ad, Bd


def discrete_euler(v, th, dt):
    """HC/models/linearization.py:98-118 via :160-188 method='euler'."""
    A, B = jacobians(v, th)
    return np.eye(3) + A * dt, B * dt


def discrete_exact(v, th, dt):
    """HC/models/linearization.py:120-158 (augmented expm)."""
    A, B = jacobians(v, th)
    Ad = expm(A * dt)
    aug = np.zeros((5, 5))
    aug[:3, :3] = A * dt
    aug[:3, 3:] = B * dt
    return Ad, expm(aug)[:3, 3:]


def lqr_gain(v, th, Qd, Rd, dt, guard=True):
    """HC/controllers/lqr_controller.py:117-132 (guard=True) / :232-240 (guard=False).

    Returns (K (2,3), P (3,3))."""
    if guard and abs(v) < 1e-6:
        v = 0.01
    Ad, Bd = discrete_explicit(v, th, dt)
    Q, R = np.diag(np.asarray(Qd, float)), np.diag(np.asarray(Rd, float))
    P = solve_discrete_are(Ad, Bd, Q, R)
    K = np.linalg.solve(R + Bd.T @ P @ Bd, Bd.T @ P @ Ad)
    return K, P


def wrap_angle(a):
    """HC/controllers/lqr_controller.py:244-250 (closed interval, +-2pi loops)."""
    while a > np.pi:
        a -= 2 * np.pi
    while a < -np.pi:
        a += 2 * np.pi
    return a


def lqr_control(x, x_ref, u_ref, K, v_max, w_max):
    """HC/controllers/lqr_controller.py:149-189."""
    e = np.array(x, float) - np.array(x_ref, float)
    e[2] = wrap_angle(e[2])
    u = np.array(u_ref, float) - K @ e
    return np.array([np.clip(u[0], -v_max, v_max), np.clip(u[1], -w_max, w_max)]), e


def plant_step(x, u, dt, v_max, w_max):
    """HC/models/differential_drive.py:138-172 (clip -> Euler -> wrap)."""
    v = float(np.clip(u[0], -v_max, v_max))
    w = float(np.clip(u[1], -w_max, w_max))
    xn = np.array([x[0] + v * np.cos(x[2]) * dt, x[1] + v * np.sin(x[2]) * dt, x[2] + w * dt])
    xn[2] = wrap_angle(xn[2])
    return xn


def figure8(duration, dt=0.02, A=2.0, a=0.5):
    """HC/trajectory/reference_generator.py:86-230 -> (n,6) [t,px,py,th,v,w]."""
    t = np.arange(0, duration, dt)

    def vel(tt):
        return a * A * np.cos(a * tt), a * A * (np.cos(a * tt) ** 2 - np.sin(a * tt) ** 2)

    out = np.zeros((len(t), 6))
    for k, tk in enumerate(t):
        dpx, dpy = vel(tk)
        th = np.arctan2(dpy, dpx)
        d2x, d2y = vel(tk + dt)
        dth = wrap_angle(np.arctan2(d2y, d2x) - th)
        out[k] = [tk, A * np.sin(a * tk), A * np.sin(a * tk) * np.cos(a * tk), th,
                  np.sqrt(dpx ** 2 + dpy ** 2), dth / dt]
    return out


def segment(traj, k0, horizon):
    """HC/trajectory/reference_generator.py:299-326 (clamped at the end)."""
    idx = np.minimum(k0 + np.arange(horizon), len(traj) - 1)
    return traj[idx, 1:4].copy(), traj[idx, 4:6].copy()
