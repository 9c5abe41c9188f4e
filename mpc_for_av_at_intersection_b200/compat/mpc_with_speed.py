"""Drop-in for the reference's `lib.mpc_with_speed` (main/lib/mpc_with_speed.py), the controller used by
main/scenarios/mpc_intersection_new_ref.py: the path is never cut; instead the speed reference `cv` drops to zero
from `cutoff_idx` on and a speed-tracking weight (Q_v_yaw[0] = 20) brings the car to rest there.

    from mpc_for_av_at_intersection_b200.compat.mpc_with_speed import MPC, MAX_ACCEL

Same members as compat.mpc.MPC except (as in the reference) `MPC(cx, cy, cv, cyaw, dl, car_dimensions, dt)` and
`set_trajectory_fromarray(trajectory, cutoff_idx=999)`.  The module constants are the ones that file hard-codes
(:16-36) instead of reading the JSON.  The device kernel holds `cv` as (ref_speed, cutoff index), which covers every
`cv` the reference ever builds (:280-282: a constant with a zero tail); any other `cv` array raises."""
import numpy as np

from ..config import MpcConfig
from . import mpc as _base
from .mpc import MPCSolutionNotFoundException, smooth_yaw  # noqa: F401

NX = 4
NU = 2
T = 13
R = np.diag([0.01, 0.01])
Rd = np.diag([0.01, 1.0])
Q_v_yaw = np.diag([20, 0.5])
Qf = np.diag([1.0, 1.0, 0., 0.5]) * T
GOAL_DIS = 1.5
STOP_SPEED = 0.5 / 3.6
MAX_TIME = 13.0
MAX_ITER = 1
DU_TH = 0.1
MAX_DSTEER = np.deg2rad(30.0)
MAX_ACCEL = 2.0
MAX_DECEL = -5
MAX_SPEED = 25 / 3.6


def _split_cv(cv):
    """cv -> (ref_speed, cutoff index): a constant, optionally followed by a zero tail."""
    cv = np.asarray(cv, float)
    nz = np.flatnonzero(cv != 0.0)
    if len(nz) == 0:
        return 0.0, 0
    cut = int(nz[-1]) + 1
    if not np.all(cv[:cut] == cv[0]):
        raise NotImplementedError("cv must be constant up to a zero tail (all the reference ever builds, mpc_with_speed.py:280-282)")
    return float(cv[0]), (cut if cut < len(cv) else 999)


class MPC(_base.MPC):
    def __init__(self, cx, cy, cv, cyaw, dl, car_dimensions, dt=0.2):
        self.cv = cv
        self._ref_speed, self._cutoff = _split_cv(cv)
        super().__init__(cx, cy, cyaw, dl, car_dimensions, dt)

    def _base_config(self) -> MpcConfig:
        return MpcConfig.with_speed(ref_speed=self._ref_speed)

    def _after_new_session(self):
        self._ses.set_speed_cutoff([self._cutoff])

    def set_trajectory_fromarray(self, trajectory: np.ndarray, cutoff_idx: int = 999):
        """main/lib/mpc_with_speed.py:276-282."""
        super().set_trajectory_fromarray(trajectory)
        self.cv = np.full_like(self.cyaw, MAX_SPEED)
        if cutoff_idx != 999:
            self.cv[cutoff_idx:] = 0
        if self._ref_speed != MAX_SPEED:                  # the constructor's cv had another level: new session config
            self._ref_speed = MAX_SPEED
            self._cfg.ref_speed = MAX_SPEED
            self._sync_state_to_session()
            self._cutoff = int(cutoff_idx)
            self._upload(self._full)
            return
        if int(cutoff_idx) != self._cutoff:
            self._cutoff = int(cutoff_idx)
            self._ses.set_speed_cutoff([self._cutoff])
