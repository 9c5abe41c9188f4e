"""BatchedMPC: B independent MPC.step calls per launch on one B200.

Host-side mirror of the reference's `MPC` object state (main/lib/mpc.py:243-299) as [B,...] CUDA
tensors; torch only owns the memory and the stream, all arithmetic is in libmpcb.so."""
import ctypes as C
from typing import Optional, Sequence

import numpy as np
import torch

from . import _lib
from .config import MpcConfig
from .courses import CourseSet  # noqa: F401

ST_OPTIMAL, ST_INACCURATE, ST_FAILED, ST_NEAREST = 0, 1, 2, 3


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


class BatchedMPC:
    def __init__(self, cfg: MpcConfig, courses: CourseSet, batch: int, device: int = 0):
        if not torch.cuda.is_available():
            raise _lib.MpcbError("BatchedMPC needs a CUDA device (no CPU fallback)")
        self.cfg, self.courses, self.B, self.device_index = cfg, courses, int(batch), int(device)
        self.dev = torch.device("cuda", device)
        self._lib = _lib.lib()
        self._h = C.c_void_p()
        ccfg = cfg.to_c()
        _lib.check(self._lib.mpcb_create(C.byref(ccfg), self.B, device, C.byref(self._h)), None, "mpcb_create")
        _lib.check(self._lib.mpcb_set_courses_host(
            self._h, courses.packed.ctypes.data_as(C.c_void_p), courses.length.ctypes.data_as(C.c_void_p),
            courses.dl.ctypes.data_as(C.c_void_p), courses.R, courses.ncmax), self._h, "mpcb_set_courses_host")
        T, B, f64, i32 = cfg.T, self.B, torch.float64, torch.int32
        z = lambda *s, dtype=f64: torch.zeros(*s, dtype=dtype, device=self.dev)
        self.route = z(B, dtype=i32)
        self.ncourse = z(B, dtype=i32)
        self.L = None
        self.target_ind = z(B, dtype=i32)
        self.warm = z(B, dtype=i32)
        self.oa, self.od = z(B, T), z(B, T)
        self.traj, self.xref = z(B, 4, T + 1), z(B, 4, T + 1)
        self.di_ai = z(B, 2)
        self.xref_deviation = z(B)
        self.status, self.iters = z(B, dtype=i32), z(B, dtype=i32)
        self.goal = z(B, 2)
        self._goal_out = z(B, dtype=i32)

    # ---- configuration of the batch --------------------------------------------------------------
    def set_routes(self, route, ncourse=None, L=None):
        """route[B]: course id per problem; ncourse[B]: cut length (default: full); L[B]: wheelbase."""
        route = torch.as_tensor(route, dtype=torch.int32)
        self.route.copy_(route)
        full = torch.as_tensor(self.courses.length)[route.long()]
        self.ncourse.copy_(full if ncourse is None else torch.as_tensor(ncourse, dtype=torch.int32))
        self.goal.copy_(torch.as_tensor(self.courses.goal)[route.long()])
        if L is not None:
            self.L = torch.as_tensor(L, dtype=torch.float64).to(self.dev).contiguous()

    def set_course_lengths(self, ncourse):
        """The only thing set_trajectory_fromarray ever changes is the cut length (mpc.py:275-278)."""
        self.ncourse.copy_(torch.as_tensor(ncourse, dtype=torch.int32))

    def reset(self):
        for t in (self.target_ind, self.warm, self.oa, self.od, self.di_ai):
            t.zero_()

    # ---- the hot call ------------------------------------------------------------------------------
    def step(self, state: torch.Tensor, stream: Optional[torch.cuda.Stream] = None):
        """state [B,4] (x, y, yaw, v) float64 CUDA.  Returns (delta[B], accel[B]) views of di_ai."""
        assert state.is_cuda and state.dtype == torch.float64 and state.shape == (self.B, 4) and state.is_contiguous()
        st = stream or torch.cuda.current_stream(self.dev)
        _lib.check(self._lib.mpcb_step(
            self._h, self.B, _ptr(state), _ptr(self.route), _ptr(self.ncourse), _ptr(self.L), _ptr(self.target_ind),
            _ptr(self.warm), _ptr(self.oa), _ptr(self.od), _ptr(self.traj), _ptr(self.xref), _ptr(self.di_ai),
            _ptr(self.xref_deviation), _ptr(self.status), _ptr(self.iters), C.c_void_p(st.cuda_stream)),
            self._h, "mpcb_step")
        return self.di_ai[:, 0], self.di_ai[:, 1]

    def plant_step(self, state: torch.Tensor, active: Optional[torch.Tensor] = None, stream=None):
        """Simulation.step for the batch, in place on `state` (main/lib/simulation.py:35-47)."""
        st = stream or torch.cuda.current_stream(self.dev)
        _lib.check(self._lib.mpcb_plant_step(self._h, self.B, _ptr(state), _ptr(self.di_ai), _ptr(self.L),
                                             _ptr(active), C.c_void_p(st.cuda_stream)), self._h, "mpcb_plant_step")
        return state

    def is_goal(self, state: torch.Tensor, stream=None) -> torch.Tensor:
        st = stream or torch.cuda.current_stream(self.dev)
        _lib.check(self._lib.mpcb_is_goal(self._h, self.B, _ptr(state), _ptr(self.goal), _ptr(self.target_ind),
                                          _ptr(self.ncourse), _ptr(self._goal_out), C.c_void_p(st.cuda_stream)),
                   self._h, "mpcb_is_goal")
        return self._goal_out

    # ---- host-buffer path (pinned staging inside the library) --------------------------------------------
    def step_host(self, state, route, ncourse, target_ind, warm, oa, od, di_ai, L=None, want_traj=True):
        """numpy in / numpy out through mpcb_step_host; in/out arrays are updated in place."""
        B, T = len(state), self.cfg.T
        out = dict(status=np.zeros(B, np.int32), iters=np.zeros(B, np.int32), xdev=np.zeros(B))
        if want_traj:
            out["traj"] = np.zeros((B, 4, T + 1))
            out["xref"] = np.zeros((B, 4, T + 1))
        p = lambda a: a.ctypes.data_as(C.c_void_p) if a is not None else None
        for a, dt in ((state, np.float64), (route, np.int32), (ncourse, np.int32), (target_ind, np.int32),
                      (warm, np.int32), (oa, np.float64), (od, np.float64), (di_ai, np.float64)):
            assert a.dtype == dt and a.flags.c_contiguous
        _lib.check(self._lib.mpcb_step_host(
            self._h, B, p(state), p(route), p(ncourse), p(L), p(target_ind), p(warm), p(oa), p(od),
            p(out.get("traj")), p(out.get("xref")), p(di_ai), p(out["xdev"]), p(out["status"]), p(out["iters"])),
            self._h, "mpcb_step_host")
        return out

    @property
    def launches(self):
        return int(self._lib.mpcb_launch_count(self._h))

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._lib.mpcb_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
