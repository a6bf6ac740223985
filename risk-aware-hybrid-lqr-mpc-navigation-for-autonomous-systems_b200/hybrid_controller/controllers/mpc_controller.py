"""Drop-in ``MPCController`` backed by the batched MPC QP kernel (K3).

Mirrors /root/reference/src/hybrid_controller/hybrid_controller/controllers/mpc_controller.py
(:33-571): ``Obstacle``, ``MPCSolution``, constructor signature/defaults, public attributes,
``solve`` (LTI, absolute state, reference padding), ``solve_with_ltv`` (LTV error state, move
blocking, theta unwrap, cold-start ramp), ``_get_fallback_solution``, ``get_warm_start``,
``reset``, ``set_obstacles``.  CVXPY -> OSQP is replaced by libb2c.so; the ``solver`` string is
accepted and ignored; there is no CPU fallback.  ``*_batch`` methods are additions.
"""
import time
from dataclasses import dataclass
from typing import Any, Dict, List, Optional

import numpy as np

from .. import _native as nat
from ..models.linearization import Linearizer


@dataclass
class Obstacle:
    """Circular obstacle (mpc_controller.py:33-46)."""
    x: float
    y: float
    radius: float

    def distance_to(self, px: float, py: float) -> float:
        return np.sqrt((px - self.x)**2 + (py - self.y)**2)

    def is_collision(self, px: float, py: float, d_safe: float) -> bool:
        return self.distance_to(px, py) < self.radius + d_safe


@dataclass
class MPCSolution:
    """mpc_controller.py:49-59."""
    status: str
    optimal_control: np.ndarray
    control_sequence: np.ndarray
    predicted_states: np.ndarray
    cost: float
    solve_time_ms: float
    slack_used: bool
    iterations: int


class MPCController:
    def __init__(self, horizon: int = 10,
                 Q_diag: list = None, R_diag: list = None, P_diag: list = None,
                 d_safe: float = 0.3, slack_penalty: float = 5000.0,
                 v_max: float = 1.0, omega_max: float = 1.5,
                 dt: float = 0.02, solver: str = "OSQP",
                 block_size: int = 1):
        # mpc_controller.py:111-148
        self.N = horizon
        self.dt = dt
        self.d_safe = d_safe
        self.slack_penalty = slack_penalty
        self.v_max = v_max
        self.omega_max = omega_max
        self.solver = solver
        self.block_size = block_size
        self.N_blocks = (horizon + block_size - 1) // block_size
        if Q_diag is None:
            Q_diag = [10.0, 10.0, 50.0]
        if R_diag is None:
            R_diag = [0.1, 0.1]
        if P_diag is None:
            P_diag = [20.0, 20.0, 40.0]
        self.Q = np.diag(Q_diag)
        self.R = np.diag(R_diag)
        self.P = np.diag(P_diag)
        self.linearizer = Linearizer(dt=dt)
        self._prev_solution: Optional[np.ndarray] = None
        self._prev_states: Optional[np.ndarray] = None
        self._step_count = 0
        self._ramp_up_steps = 10
        self.nx = 3
        self.nu = 2
        self.engine = 0      # addition: K3 kernel choice (0 auto, 1 general, 2 slot kernel), see include/b2c.h

    # -- engine plumbing -------------------------------------------------------------------------
    @property
    def _ctx(self):
        return nat.default_context(-1)

    def _params(self, kind, M, soft=True):
        for W in (self.Q, self.R, self.P):
            if np.any(np.asarray(W) != np.diag(np.diag(W))):
                raise NotImplementedError("the MPC kernel supports diagonal Q, R, P")
        return nat.mpc_params(kind, self.N, M, self.block_size if kind == nat.MPC_LTV else 1,
                              np.diag(self.Q), np.diag(self.R), np.diag(self.P), self.d_safe,
                              self.slack_penalty, self.v_max, self.omega_max, self.dt, soft=soft,
                              ramp_up_steps=self._ramp_up_steps, engine=self.engine)

    @staticmethod
    def _obs_array(obstacles):
        if not obstacles:
            return np.zeros((0, 3))
        return np.array([[o.x, o.y, o.radius] for o in obstacles], dtype=float)

    def _solution(self, out, b, solve_time) -> MPCSolution:
        ok = out["status"][b] == nat.ST_OK
        return MPCSolution(
            status="optimal" if ok else "fallback",
            optimal_control=out["u0"][b].copy(),
            control_sequence=out["useq"][b].copy(),
            predicted_states=out["xpred"][b].copy(),
            cost=float(out["cost"][b]),
            solve_time_ms=solve_time,
            slack_used=bool(out["slack_used"][b]) if ok else False,
            iterations=0)                       # the reference always reports 0 (:310, :519)

    # -- reference API ---------------------------------------------------------------------------
    def solve(self, x0: np.ndarray, x_refs: np.ndarray, u_refs: np.ndarray,
              obstacles: List[Obstacle] = None,
              use_soft_constraints: bool = True) -> MPCSolution:
        """LTI absolute-state MPC (mpc_controller.py:150-314)."""
        start_time = time.perf_counter()
        if obstacles is None:
            obstacles = []
        x_refs = np.asarray(x_refs, dtype=float)
        u_refs = np.asarray(u_refs, dtype=float)
        if x_refs.shape[0] < self.N + 1:        # :172-177 pad with last reference
            pad = np.zeros((self.N + 1, 3))
            pad[:x_refs.shape[0]] = x_refs
            pad[x_refs.shape[0]:] = x_refs[-1]
            x_refs = pad
        if u_refs.shape[0] < self.N:            # :179-183
            pad = np.zeros((self.N, 2))
            pad[:u_refs.shape[0]] = u_refs
            pad[u_refs.shape[0]:] = u_refs[-1]
            u_refs = pad
        obs = self._obs_array(obstacles)
        prm = self._params(nat.MPC_LTI, len(obs), soft=use_soft_constraints)
        out = nat.mpc_host(self._ctx, prm, np.asarray(x0, float)[None], x_refs[None, :self.N + 1],
                           u_refs[None, :self.N], obs[None], None, want_seq=True)
        sol = self._solution(out, 0, (time.perf_counter() - start_time) * 1000)
        if sol.status == "optimal":             # :299-300
            self._prev_solution = sol.control_sequence
            self._prev_states = sol.predicted_states
        return sol

    def _get_fallback_solution(self, x0: np.ndarray, x_refs: np.ndarray,
                               u_refs: np.ndarray, solve_time: float) -> MPCSolution:
        """P-control fallback (mpc_controller.py:316-343); the kernel applies the same formula to
        every non-converged instance, this host copy serves direct callers."""
        error = np.asarray(x0, float) - np.asarray(x_refs[0], float)
        error[2] = self._normalize_angle(error[2])
        K_p = np.array([[1.0, 0.0, 0.0],
                        [0.0, 0.0, 0.5]])
        u_fallback = self._clip_control(u_refs[0] + (-K_p @ error))
        return MPCSolution(status="fallback", optimal_control=u_fallback,
                           control_sequence=np.tile(u_fallback, (self.N, 1)),
                           predicted_states=np.tile(x0, (self.N + 1, 1)),
                           cost=float('inf'), solve_time_ms=solve_time, slack_used=False,
                           iterations=0)

    def solve_with_ltv(self, x0: np.ndarray, x_refs: np.ndarray,
                       u_refs: np.ndarray, obstacles: List[Obstacle] = None,
                       use_soft_constraints: bool = True) -> MPCSolution:
        """LTV error-state MPC with move blocking (mpc_controller.py:345-522)."""
        start_time = time.perf_counter()
        if obstacles is None:
            obstacles = []
        x_refs = np.asarray(x_refs, dtype=float)
        u_refs = np.asarray(u_refs, dtype=float)
        if x_refs.shape[0] < self.N + 1 or u_refs.shape[0] < self.N:
            # the reference indexes x_refs[k], k <= N and u_refs[k], k < N without padding (Q6)
            raise IndexError(f"solve_with_ltv needs {self.N + 1} reference states and {self.N} "
                             f"reference controls, got {x_refs.shape[0]} / {u_refs.shape[0]}")
        obs = self._obs_array(obstacles)
        prm = self._params(nat.MPC_LTV, len(obs), soft=use_soft_constraints)
        sc = np.array([self._step_count], dtype=np.int32)
        out = nat.mpc_host(self._ctx, prm, np.asarray(x0, float)[None], x_refs[None, :self.N + 1],
                           u_refs[None, :self.N], obs[None], sc, want_seq=True)
        sol = self._solution(out, 0, (time.perf_counter() - start_time) * 1000)
        if sol.status == "optimal":             # :506-509
            self._step_count = int(out["step_count"][0])
            self._prev_solution = sol.control_sequence
            self._prev_states = sol.predicted_states
        return sol

    def get_warm_start(self) -> Optional[np.ndarray]:
        """mpc_controller.py:524-538."""
        if self._prev_solution is None:
            return None
        warm_start = np.zeros_like(self._prev_solution)
        warm_start[:-1] = self._prev_solution[1:]
        warm_start[-1] = self._prev_solution[-1]
        return warm_start

    def _normalize_angle(self, angle: float) -> float:
        while angle > np.pi:
            angle -= 2 * np.pi
        while angle < -np.pi:
            angle += 2 * np.pi
        return angle

    def reset(self):
        """mpc_controller.py:548-552."""
        self._step_count = 0
        self._prev_solution = None
        self._prev_states = None

    def _clip_control(self, u: np.ndarray) -> np.ndarray:
        return np.array([
            np.clip(u[0], -self.v_max, self.v_max),
            np.clip(u[1], -self.omega_max, self.omega_max)
        ])

    def set_obstacles(self, obstacles: List[Dict[str, float]]) -> List[Obstacle]:
        """mpc_controller.py:561-571."""
        return [Obstacle(x=o['x'], y=o['y'], radius=o['radius']) for o in obstacles]

    # -- batch additions ---------------------------------------------------------------------------
    def solve_with_ltv_batch(self, x0, x_refs, u_refs, obstacles, step_count=None,
                             use_soft_constraints: bool = True, want_seq: bool = True) -> Dict[str, Any]:
        """B independent solve_with_ltv instances in one launch.

        x0 (B,3), x_refs (B,N+1,3), u_refs (B,N,2), obstacles (B,M,3) as [x,y,radius] (or (M,3),
        broadcast), step_count (B,) int32 per-instance successful-solve counters (None = ramp off).
        Returns dict: u0 (B,2), cost (B,), status (B,) [0 = optimal, else fallback values], iters,
        slack_used, useq (B,N,2), xpred (B,N+1,3), step_count (updated)."""
        return self._batch(nat.MPC_LTV, x0, x_refs, u_refs, obstacles, step_count,
                           use_soft_constraints, want_seq)

    def solve_batch(self, x0, x_refs, u_refs, obstacles, use_soft_constraints: bool = True,
                    want_seq: bool = True) -> Dict[str, Any]:
        """B independent solve() instances (references already N+1 / N rows)."""
        return self._batch(nat.MPC_LTI, x0, x_refs, u_refs, obstacles, None,
                           use_soft_constraints, want_seq)

    def solve_with_ltv_table_batch(self, x0, table, k0, obstacles, step_count=None, want_seq: bool = True) -> Dict[str, Any]:
        """solve_with_ltv for B instances that share one reference table: instance i solves on
        ReferenceTrajectoryGenerator.get_trajectory_segment(k0[i], N+1) of `table` (rows,6), gathered on the device
        (what run_simulation.py:247 does per robot before each call).  x0 (B,3), k0 (B,), obstacles (B,M,3) or (M,3).
        Same outputs as solve_with_ltv_batch, bit-identical to it on the materialised segments."""
        x0 = np.ascontiguousarray(x0, float).reshape(-1, 3)
        B = x0.shape[0]
        obstacles = np.asarray(obstacles, float)
        if obstacles.ndim == 2:
            obstacles = np.broadcast_to(obstacles[None], (B,) + obstacles.shape)
        M = obstacles.shape[1] if obstacles.size else 0
        prm = self._params(nat.MPC_LTV, M, soft=True)
        return nat.mpc_table_host(self._ctx, prm, table, k0, x0, obstacles, step_count, want_seq=want_seq)

    def _batch(self, kind, x0, x_refs, u_refs, obstacles, step_count, soft, want_seq):
        x0 = np.ascontiguousarray(x0, float).reshape(-1, 3)
        B = x0.shape[0]
        obstacles = np.asarray(obstacles, float)
        if obstacles.ndim == 2:
            obstacles = np.broadcast_to(obstacles[None], (B,) + obstacles.shape)
        M = obstacles.shape[1] if obstacles.size else 0
        prm = self._params(kind, M, soft=soft)
        t0 = time.perf_counter()
        out = nat.mpc_host(self._ctx, prm, x0, x_refs, u_refs, obstacles, step_count, want_seq=want_seq)
        out["solve_time_ms"] = (time.perf_counter() - t0) * 1000
        return out
