"""Host mirror of the reference's risk supervisor (HC/controllers/risk_metrics.py:21-222): same class, method and
field names.  Scalar calls are plain numpy (there is nothing to batch in one assessment); the batched form runs on
the device inside ``hybrid_controller.rollout`` (csrc/rollout.cu: roll_risk)."""
from dataclasses import dataclass
from typing import Dict, List, Optional, Tuple

import numpy as np


@dataclass
class RiskAssessment:
    distance_risk: float
    predictive_risk: float
    combined_risk: float
    min_obstacle_distance: float
    nearest_obstacle_id: int
    use_mpc: bool
    risk_level: str


class RiskMetrics:
    def __init__(self, d_safe: float = 0.3, d_trigger: float = 1.0, alpha: float = 0.6, beta: float = 0.4,
                 threshold_low: float = 0.2, threshold_medium: float = 0.5, threshold_high: float = 0.8):
        self.d_safe, self.d_trigger = d_safe, d_trigger
        self.threshold_low, self.threshold_medium, self.threshold_high = threshold_low, threshold_medium, threshold_high
        total = alpha + beta                         # weights are normalised to sum to one, :76-79
        self.alpha, self.beta = alpha / total, beta / total

    def compute_distance_risk(self, state: np.ndarray, obstacles: List[Dict]) -> Tuple[float, float, int]:
        """max over obstacles of the ramp 1 -> 0 between d_safe and d_trigger of the edge distance, :84-132."""
        if not obstacles:
            return 0.0, float('inf'), -1
        best, nearest, worst = float('inf'), -1, 0.0
        for i, o in enumerate(obstacles):
            edge = np.sqrt((state[0] - o['x']) ** 2 + (state[1] - o['y']) ** 2) - o['radius']
            if edge < best:
                best, nearest = edge, i
            if edge <= self.d_safe:
                r = 1.0
            elif edge >= self.d_trigger:
                r = 0.0
            else:
                r = 1.0 - (edge - self.d_safe) / (self.d_trigger - self.d_safe)
            worst = max(worst, r)
        return worst, best, nearest

    def compute_predictive_risk(self, predicted_states: Optional[np.ndarray], obstacles: List[Dict]) -> float:
        """time-weighted severity of predicted safety-zone violations, scaled by 5 / (N * n_obstacles), :134-176."""
        if not obstacles or predicted_states is None or len(predicted_states) == 0:
            return 0.0
        n = len(predicted_states)
        total = 0.0
        for k, s in enumerate(predicted_states):
            for o in obstacles:
                edge = np.sqrt((s[0] - o['x']) ** 2 + (s[1] - o['y']) ** 2) - o['radius']
                if edge < self.d_safe:
                    total += (1.0 - (k / n) * 0.5) * ((self.d_safe - edge) / self.d_safe)
        return min(1.0, total / (n * len(obstacles)) * 5)

    def assess_risk(self, state: np.ndarray, obstacles: List[Dict],
                    predicted_states: Optional[np.ndarray] = None) -> RiskAssessment:
        d, dmin, nearest = self.compute_distance_risk(state, obstacles)
        p = self.compute_predictive_risk(predicted_states, obstacles) if predicted_states is not None else 0.0
        c = self.alpha * d + self.beta * p
        level = ("low" if c < self.threshold_low else "medium" if c < self.threshold_medium
                 else "high" if c < self.threshold_high else "critical")
        return RiskAssessment(distance_risk=d, predictive_risk=p, combined_risk=c, min_obstacle_distance=dmin,
                              nearest_obstacle_id=nearest, use_mpc=c >= self.threshold_low, risk_level=level)

    def get_risk_summary(self, a: RiskAssessment) -> str:
        return (f"Risk: {a.risk_level.upper()} (combined={a.combined_risk:.2f}, dist={a.distance_risk:.2f}, "
                f"pred={a.predictive_risk:.2f}, min_d={a.min_obstacle_distance:.2f}m)")
