"""Drop-in for the reference's `lib.mpc_jerk` (main/lib/mpc_jerk.py): the same `MPC` class members as `lib.mpc`, the NX = 5
model (the acceleration input is also integrated into a fifth state with a free initial value that feeds the speed,
:61-86, and its increments are penalised, :190) and the constants that file hard-codes instead of reading the JSON
(:16-39,167-171: T 13, w_perp 10, w_para 1, Rd (.3, 1), MAX_DECEL -5, STOP_SPEED .5/3.6).

    from mpc_for_av_at_intersection_b200.compat.mpc_jerk import MPC, MAX_ACCEL      # instead of `from lib.mpc_jerk import ...`

A B=1 view of the batched engine like compat.mpc; the solve runs in the jerk instantiation of the CTA-per-problem kernel
(`mpcb_config.jerk_model`).  No script of the reference imports lib.mpc_jerk; the goldens come from the module itself
(oracle/gen_golden_jerk.py)."""
import numpy as np

from ..config import MpcConfig
from . import mpc as _base
from .mpc import MPCSolutionNotFoundException, smooth_yaw  # noqa: F401

NX = 5
NU = 2
T = 13
R = np.diag([0.01, 0.01])
Rd = np.diag([.3, 1.0])
Q_v_yaw = np.diag([0., 0.5])
Qf = np.diag([1.0, 1.0, 0., 0.5, 0]) * T
GOAL_DIS = 1.5
STOP_SPEED = 0.5 / 3.6
MAX_TIME = 13.0
jerk_penalty_weight = 1
MAX_ITER = 1
DU_TH = 0.1
MAX_DSTEER = np.deg2rad(30.0)
MAX_ACCEL = 2.0
MAX_DECEL = -5


class MPC(_base.MPC):
    @staticmethod
    def _base_config() -> MpcConfig:
        return MpcConfig.jerk(jerk_weight=float(jerk_penalty_weight))
