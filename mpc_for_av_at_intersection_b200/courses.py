"""CourseSet: the A* paths (cx, cy, cyaw, dl of MPC.__init__, main/lib/mpc.py:243-263) shared by a batch."""
import ctypes as C
import os
from typing import Optional, Sequence

import numpy as np

from . import _lib

_ROUTES_NPZ = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "intersection_routes.npz")


def smooth_yaw_host(yaw):
    """smooth_yaw (main/lib/mpc.py:43-55) in place, without the native library; bit-identical to mpcb_smooth_yaw
    (tests/test_abi_exports.py)."""
    for i in range(len(yaw) - 1):
        dyaw = yaw[i + 1] - yaw[i]
        while dyaw >= np.pi / 2.0:
            yaw[i + 1] -= np.pi * 2.0
            dyaw = yaw[i + 1] - yaw[i]
        while dyaw <= -np.pi / 2.0:
            yaw[i + 1] += np.pi * 2.0
            dyaw = yaw[i + 1] - yaw[i]
    return yaw


class CourseSet:
    """Routes shared by the batch: [R,3,ncmax] (x | y | smoothed yaw), lengths and tick dl."""

    def __init__(self, courses: Sequence[np.ndarray], dl: Optional[Sequence[float]] = None, smooth=True):
        self.R = len(courses)
        self.length = np.array([len(c) for c in courses], dtype=np.int32)
        self.ncmax = int(self.length.max())
        self.packed = np.zeros((self.R, 3, self.ncmax), dtype=np.float64)
        lib = _lib.lib() if smooth and smooth != "numpy" else None
        for r, c in enumerate(courses):
            c = np.ascontiguousarray(np.asarray(c, dtype=np.float64)[:, :3])
            yaw = np.ascontiguousarray(c[:, 2])
            if smooth == "numpy":      # same scan in plain Python: for processes that must not map libmpcb.so (bench.py's CPU arm)
                smooth_yaw_host(yaw)
            elif smooth:       # MPC.__init__ smooths cyaw once (main/lib/mpc.py:257)
                lib.mpcb_smooth_yaw(yaw.ctypes.data_as(C.c_void_p), len(yaw))
            n = len(c)
            self.packed[r, 0, :n], self.packed[r, 1, :n], self.packed[r, 2, :n] = c[:, 0], c[:, 1], yaw
            self.packed[r, :, n:] = self.packed[r, :, n - 1:n]
        if dl is None:   # dl = |p0 - p1| as main/scenarios/mpc_intersection.py:69
            dl = [float(np.linalg.norm(np.asarray(c)[0, :2] - np.asarray(c)[1, :2])) for c in courses]
        self.dl = np.asarray(dl, dtype=np.float64)
        self.goal = np.array([[self.packed[r, 0, n - 1], self.packed[r, 1, n - 1]] for r, n in enumerate(self.length)])


def load_intersection_routes():
    """The 12 (start_pos, turn_indicator) routes of main/envs/intersection.py planned by the reference's
    MotionPrimitiveSearch (recorded by oracle/gen_golden.py; raw yaw, before smooth_yaw)."""
    d = np.load(_ROUTES_NPZ)
    return [d[f"route_{i}"].copy() for i in range(len(d["length"]))]
