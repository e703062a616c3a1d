"""`CostWeights` with the reference's fields (src/mpc/cost_functions.py:39-103).  The 3-DoF hot path reads the
diagonal of the 7-state sub-block [m, r, v] and of R; the defaults reproduce osqp_rti.py:168-182
(Q = diag(0,10,10,10,1,1,1), R = 0.01 I, P = 10 Q)."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import numpy as np


@dataclass
class CostWeights:
    Q: Optional[np.ndarray] = None
    R: Optional[np.ndarray] = None
    P: Optional[np.ndarray] = None
    w_position: float = 10.0
    w_velocity: float = 1.0
    w_attitude: float = 5.0
    w_omega: float = 0.1
    w_mass: float = 0.0
    w_thrust: float = 0.01
    w_fuel: float = 0.1
    terminal_weight: float = 10.0

    def __post_init__(self):
        if self.Q is None:
            self.Q = self._build_Q_matrix()
        if self.R is None:
            self.R = self._build_R_matrix()
        if self.P is None:
            self.P = self.terminal_weight * self.Q

    def _build_Q_matrix(self):
        Q = np.zeros((14, 14))
        Q[0, 0] = self.w_mass
        Q[1:4, 1:4] = self.w_position * np.eye(3)
        Q[4:7, 4:7] = self.w_velocity * np.eye(3)
        Q[8, 8] = self.w_attitude
        Q[9, 9] = self.w_attitude
        Q[11:14, 11:14] = self.w_omega * np.eye(3)
        return Q

    def _build_R_matrix(self):
        return self.w_thrust * np.eye(3)

    # 3-DoF views used by the QP path
    def diag_3dof(self):
        """(Q_diag (7,), Qf_diag (7,), R_diag (3,)) of the [m, r, v] sub-block."""
        Q = np.asarray(self.Q)
        P = np.asarray(self.P)
        return np.diag(Q)[:7].copy(), np.diag(P)[:7].copy(), np.diag(np.asarray(self.R)).copy()
