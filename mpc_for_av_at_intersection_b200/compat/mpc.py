"""Drop-in for the reference's `lib.mpc` (main/lib/mpc.py): same module constants, same `MPC` class members,
same argument order and error behaviour -- a B=1 view onto the batched CUDA engine (MpcSession).

    from mpc_for_av_at_intersection_b200.compat.mpc import MPC, MAX_ACCEL      # instead of `from lib.mpc import ...`

Configuration: like the reference (main/lib/mpc.py:13-36) the JSON is read at import; the search order is
$MPCB_CONFIG, then '../config/mpc_config.json' relative to the cwd (the reference's own rule, so running from
main/scenarios/ picks up the reference's file), then the packaged copy of the reference defaults.

Differences a caller can observe, all deliberate:
  * the course arrays are snapshotted to the device at construction / set_trajectory_fromarray instead of
    being held as live numpy views (mutating them afterwards is not seen until the next set_trajectory call);
  * the solve runs on the GPU: without a CUDA device construction raises (there is no CPU fallback).
"""
import json
import math
import os
import sys

import numpy as np

from .. import _lib
from ..config import MpcConfig
from ..courses import CourseSet
from ..session import MpcSession


def _find_config():
    env = os.environ.get("MPCB_CONFIG")
    if env:
        return env
    rel = os.path.join("..", "config", "mpc_config.json")
    if os.path.exists(rel):
        return rel
    return os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "config", "mpc_config.json")


CONFIG_PATH = _find_config()
with open(CONFIG_PATH, "r") as _f:
    _config = json.load(_f)

# the globals of main/lib/mpc.py:17-36, same names, same derived units
NX = _config["NX"]
NU = _config["NU"]
T = _config["T"]
w_perp = _config["w_perp"]
w_para = _config["w_para"]
R = np.diag(_config["R"])
Rd = np.diag(_config["Rd"])
Q_v_yaw = np.diag(_config["Q_v_yaw"])
Qf = np.diag(_config["Qf"]) * T
GOAL_DIS = _config["GOAL_DIS"]
STOP_SPEED = _config["STOP_SPEED"]
MAX_TIME = _config["MAX_TIME"]
MAX_ITER = _config["MAX_ITER"]
DU_TH = _config["DU_TH"]
MAX_DSTEER = np.deg2rad(_config["MAX_DSTEER"])
MAX_ACCEL = _config["MAX_ACCEL"]
MAX_DECEL = _config["MAX_DECEL"]


class MPCSolutionNotFoundException(RuntimeError):
    pass


def smooth_yaw(yaw):
    """main/lib/mpc.py:43-55 -- in place (also on a strided view such as trajectory[:, 2]); returns yaw."""
    import ctypes as C
    buf = np.ascontiguousarray(yaw, dtype=np.float64)
    if buf is yaw or np.shares_memory(buf, yaw):
        _lib.lib().mpcb_smooth_yaw(buf.ctypes.data_as(C.c_void_p), len(buf))
    else:
        _lib.lib().mpcb_smooth_yaw(buf.ctypes.data_as(C.c_void_p), len(buf))
        yaw[:] = buf
    return yaw


class MPC:
    def __init__(self, cx, cy, cyaw, dl, car_dimensions, dt=0.2):
        self.car_dimensions = car_dimensions
        self.dl = dl
        self.dt = dt
        self.cx = cx
        self.cy = cy
        self.cyaw = smooth_yaw(cyaw)                   # in place on the caller's array, mpc.py:257
        self.goal = [cx[-1], cy[-1]]
        self.target_ind = 0
        self.oa, self.odelta = None, None
        self.ox = self.oy = self.oyaw = self.ov = None
        self.xref = None
        self.di, self.ai = 0.0, 0.0
        self._cfg = self._base_config()
        self._cfg.dt = float(dt)
        self._cfg.wheelbase = float(car_dimensions.distance_back_to_front_wheel)
        self._ses = None
        self._upload(np.stack([np.asarray(cx, float), np.asarray(cy, float), np.asarray(cyaw, float)], axis=1))

    @staticmethod
    def _base_config() -> MpcConfig:
        """The module-level constants as a config object (the variants in compat/ override this)."""
        return MpcConfig.from_dict({k: v for k, v in _config.items()})

    def _after_new_session(self):
        pass

    # ---- course handling ------------------------------------------------------------------------------------
    def _upload(self, course):
        """(Re)create the B=1 session on a new full course; solver state (target_ind, warm start) carries over."""
        keep = None
        if self._ses is not None:
            keep = {k: self._ses.fetch(k) for k in ("oa", "odelta", "warm", "di_ai")}
            self._ses.close()
        self._full = np.ascontiguousarray(course[:, :3], dtype=np.float64)
        cs = CourseSet([self._full], dl=[float(self.dl)], smooth=False)
        self._ses = MpcSession(self._cfg, cs, 1).begin([0], target_ind=[self.target_ind])
        if keep is not None:
            for k, v in keep.items():
                self._ses.store(k, v)
        self._ncourse = len(self._full)
        self._after_new_session()

    def set_trajectory_fromarray(self, trajectory: np.ndarray):
        """main/lib/mpc.py:275-278: rebinds cx, cy, cyaw; target_ind, goal and the warm start are kept."""
        self.cx = trajectory[:, 0]
        self.cy = trajectory[:, 1]
        self.cyaw = trajectory[:, 2]
        n = len(trajectory)
        if 1 <= n <= len(self._full) and np.array_equal(trajectory[:, :3], self._full[:n]):
            if n != self._ncourse:                       # the usual case: a prefix cut of the planned path
                self._ses.set_course_lengths([n])
                self._ncourse = n
        else:
            self._sync_state_to_session()
            self._upload(np.asarray(trajectory, float))

    def _sync_state_to_session(self):
        self._ses.store("target_ind", [self.target_ind])

    # ---- MPC.step (main/lib/mpc.py:280-299) -----------------------------------------------------------------
    def step(self, state):
        ses = self._ses
        st = np.array([[state.x, state.y, state.yaw, state.v]], dtype=np.float64)
        # one call, one synchronisation: target_ind in (callers may set mpc.target_ind directly), `oa is None` as the warm
        # flag, the step, and every attribute the reference's callers read afterwards out
        out = ses.step_all(st, target_ind=[self.target_ind], warm=[0] if (self.oa is None or self.odelta is None) else None)
        status = int(out["status"][0])
        if status == 3:                                  # main/lib/trajectories.py:120
            raise Exception("something wrong")
        self.target_ind = int(out["target_ind"][0])
        self.xref = out["xref"][0]
        if status <= 1:
            self.oa = out["oa"][0]
            self.odelta = out["odelta"][0]
            traj = out["traj"][0]
            self.ox, self.oy, self.ov, self.oyaw = traj[0], traj[1], traj[2], traj[3]
            self.di, self.ai = float(self.odelta[0]), float(self.oa[0])
        else:
            print("Error: Cannot solve mpc..", file=sys.stderr)      # mpc.py:204-206
            self.oa = self.odelta = self.ox = self.oy = self.oyaw = self.ov = None
            self.ai = self._cfg.MAX_DECEL                             # mpc.py:294-297: emergency braking, di kept
        self.status = status
        self.iters = int(out["iters"][0])
        return self.di, self.ai

    def get_current_xref_deviation(self):
        """main/lib/mpc.py:301-308 (element-wise product as written there; raises like the reference when the
        last solve failed and ox is None)."""
        dx = self.cx[self.target_ind] - self.ox[0]
        dy = self.cy[self.target_ind] - self.oy[0]
        ang = self.cyaw[self.target_ind] + np.pi / 2
        return float(np.linalg.norm(np.array([np.cos(ang) * dx, np.sin(ang) * dy])))

    def is_goal(self, state):
        """main/lib/mpc.py:310-326."""
        d = math.hypot(state.x - self.goal[0], state.y - self.goal[1])
        isgoal = d <= self._cfg.GOAL_DIS
        if abs(self.target_ind - len(self.cx)) >= 5:
            isgoal = False
        return bool(isgoal and abs(state.v) <= self._cfg.STOP_SPEED)
