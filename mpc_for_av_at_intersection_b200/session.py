"""MpcSession: B `MPC` objects of the reference (main/lib/mpc.py:242-326) as ONE device-resident batch.

Per control step only the vehicle states cross the host boundary (`step(state[B,4]) -> (delta[B], accel[B])`,
the batched form of `di, ai = mpc.step(state)`); target_ind, the unshifted warm start, ox/oy/ov/oyaw and
xref stay in HBM inside libmpcb.so's handle and are fetched on demand, the way callers read `mpc.ox`
after a step (main/scenarios/mpc_intersection.py:297-302).  numpy in, numpy out: torch is not needed here.
There is no CPU fallback."""
import ctypes as C
from typing import Optional

import numpy as np

from . import _lib
from .config import MpcConfig

_FIELDS = {
    # name: (field id, dtype, trailing shape as a function of T)
    "state": (_lib.F_STATE, np.float64, lambda T: (4,)),
    "oa": (_lib.F_OA, np.float64, lambda T: (T,)),
    "odelta": (_lib.F_OD, np.float64, lambda T: (T,)),
    "traj": (_lib.F_TRAJ, np.float64, lambda T: (4, T + 1)),
    "xref": (_lib.F_XREF, np.float64, lambda T: (4, T + 1)),
    "di_ai": (_lib.F_DI_AI, np.float64, lambda T: (2,)),
    "xref_deviation": (_lib.F_XDEV, np.float64, lambda T: ()),
    "target_ind": (_lib.F_TARGET, np.int32, lambda T: ()),
    "warm": (_lib.F_WARM, np.int32, lambda T: ()),
    "status": (_lib.F_STATUS, np.int32, lambda T: ()),
    "iters": (_lib.F_ITERS, np.int32, lambda T: ()),
    "done": (_lib.F_DONE, np.int32, lambda T: ()),
    "ncourse": (_lib.F_NCOURSE, np.int32, lambda T: ()),
    "route": (_lib.F_ROUTE, np.int32, lambda T: ()),
    "goal": (_lib.F_GOAL, np.float64, lambda T: (2,)),
    "agent_idx": (_lib.F_AGENT_IDX, np.int32, lambda T: ()),
    "collided": (_lib.F_COLLIDED, np.int32, lambda T: ()),
    "coll_idx": (_lib.F_COLL_IDX, np.int32, lambda T: ()),
    "script": (_lib.F_SCRIPT, np.float64, lambda T: (_lib.MAX_OBSTACLES, 8)),
    "obstacles": (_lib.F_OBS, np.float64, lambda T: (_lib.MAX_OBSTACLES, 6)),
    "n_obstacles": (_lib.F_NOBS, np.int32, lambda T: ()),
    "launch_order": (_lib.F_LAUNCH_ORDER, np.int32, lambda T: ()),
}


class _DevView:
    """Zero-copy view of a session field for torch.as_tensor / cupy (CUDA array interface v3)."""

    def __init__(self, ptr, shape, dtype):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": np.dtype(dtype).str,
                                         "data": (int(ptr), False), "version": 3, "strides": None}


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class MpcSession:
    def __init__(self, cfg: MpcConfig, courses, max_batch: int, device: int = 0):
        self.cfg, self.courses, self.max_batch, self.device = cfg, courses, int(max_batch), int(device)
        self._lib = _lib.lib()
        self._h = C.c_void_p()
        ccfg = cfg.to_c()
        _lib.check(self._lib.mpcb_create(C.byref(ccfg), self.max_batch, device, C.byref(self._h)), None, "mpcb_create")
        _lib.check(self._lib.mpcb_set_courses_host(
            self._h, _p(courses.packed), _p(courses.length), _p(courses.dl), courses.R, courses.ncmax),
            self._h, "mpcb_set_courses_host")
        self.B = 0

    # ---- MPC.__init__ / set_trajectory_fromarray for the batch -----------------------------------------
    def begin(self, route, ncourse=None, L=None, target_ind=None):
        route = np.ascontiguousarray(route, dtype=np.int32)
        self.B = len(route)
        nc = None if ncourse is None else np.ascontiguousarray(ncourse, dtype=np.int32)
        Lh = None if L is None else np.ascontiguousarray(L, dtype=np.float64)
        ti = None if target_ind is None else np.ascontiguousarray(target_ind, dtype=np.int32)
        for a in (nc, Lh, ti):
            assert a is None or a.shape == (self.B,)
        _lib.check(self._lib.mpcb_session_begin(self._h, self.B, _p(route), _p(nc), _p(Lh), _p(ti)),
                   self._h, "mpcb_session_begin")
        self._di_ai = np.zeros((self.B, 2))
        self._status = np.zeros(self.B, np.int32)
        return self

    def set_course_lengths(self, ncourse):
        nc = np.ascontiguousarray(ncourse, dtype=np.int32)
        assert nc.shape == (self.B,)
        _lib.check(self._lib.mpcb_session_set_ncourse_host(self._h, _p(nc)), self._h, "mpcb_session_set_ncourse_host")

    # ---- MPC.step for the batch ---------------------------------------------------------------------------
    def step(self, state: np.ndarray):
        """state [B,4] float64 (x, y, yaw, v) on the HOST -> (delta[B], accel[B]) host views (valid until the
        next step); per-problem outcomes in self.status."""
        assert state.dtype == np.float64 and state.shape == (self.B, 4) and state.flags.c_contiguous
        _lib.check(self._lib.mpcb_session_step_host(self._h, _p(state), _p(self._di_ai), _p(self._status)),
                   self._h, "mpcb_session_step_host")
        return self._di_ai[:, 0], self._di_ai[:, 1]

    def step_all(self, state: np.ndarray, target_ind=None, warm=None):
        """MPC.step plus everything the reference's callers read afterwards, in one synchronisation
        (mpcb_session_step_all_host).  Returns dict(di_ai, status, target_ind, iters, oa, odelta, traj, xref)."""
        assert state.dtype == np.float64 and state.shape == (self.B, 4) and state.flags.c_contiguous
        B, T = self.B, self.cfg.T
        ti = None if target_ind is None else np.ascontiguousarray(target_ind, dtype=np.int32)
        wm = None if warm is None else np.ascontiguousarray(warm, dtype=np.int32)
        out = dict(di_ai=np.empty((B, 2)), status=np.empty(B, np.int32), target_ind=np.empty(B, np.int32), iters=np.empty(B, np.int32),
                   oa=np.empty((B, T)), odelta=np.empty((B, T)), traj=np.empty((B, 4, T + 1)), xref=np.empty((B, 4, T + 1)))
        _lib.check(self._lib.mpcb_session_step_all_host(
            self._h, _p(state), _p(ti), _p(wm), _p(out["di_ai"]), _p(out["status"]), _p(out["target_ind"]), _p(out["iters"]),
            _p(out["oa"]), _p(out["odelta"]), _p(out["traj"]), _p(out["xref"])), self._h, "mpcb_session_step_all_host")
        self._di_ai, self._status = out["di_ai"], out["status"]
        return out

    @property
    def status(self):
        return self._status

    def step_dev(self, cuda_stream: int = 0):
        """Step on the state already resident in the session (asynchronous on the given cudaStream_t)."""
        _lib.check(self._lib.mpcb_session_step_dev(self._h, C.c_void_p(cuda_stream)), self._h, "mpcb_session_step_dev")

    def closed_loop_dev(self, nsteps: int, history_ptr: int = 0, cuda_stream: int = 0, goal_test: bool = True):
        """nsteps x (is_goal -> done | MPC.step | Simulation.step), one launch per control step, asynchronous.
        history_ptr: device pointer to float64 [nsteps,B,8] or 0."""
        _lib.check(self._lib.mpcb_session_closed_loop_dev(
            self._h, int(nsteps), C.c_void_p(history_ptr) if history_ptr else None, 0 if goal_test else 1,
            C.c_void_p(cuda_stream)),
            self._h, "mpcb_session_closed_loop_dev")

    def multi_agent_dev(self, nsteps: int, vehicles_per_scenario: int, history_ptr: int = 0, cuda_stream: int = 0,
                        goal_test: bool = True):
        """nsteps x (is_goal | coupling against the other egos of the scenario | MPC.step | Simulation.step) on the
        resident session: problems are grouped in scenarios of `vehicles_per_scenario` consecutive egos."""
        _lib.check(self._lib.mpcb_session_multi_agent_dev(
            self._h, int(nsteps), int(vehicles_per_scenario), C.c_void_p(history_ptr) if history_ptr else None,
            0 if goal_test else 1, C.c_void_p(cuda_stream)), self._h, "mpcb_session_multi_agent_dev")

    # ---- controller variants --------------------------------------------------------------------
    def set_weights(self, weights):
        """Per-problem cost weights [B,12] (MpcConfig.weight_row order, file units) -- the batched form of
        main/lib/mpc_sensitivity.py re-reading its JSON on every solve; None = the session's config for everyone."""
        if weights is None:
            _lib.check(self._lib.mpcb_session_set_weights_host(self._h, None), self._h, "mpcb_session_set_weights_host")
            return
        w = np.ascontiguousarray(weights, dtype=np.float64)
        assert w.shape == (self.B, 12), w.shape
        _lib.check(self._lib.mpcb_session_set_weights_host(self._h, _p(w)), self._h, "mpcb_session_set_weights_host")

    def set_speed_cutoff(self, cutoff_idx):
        """mpc_with_speed.MPC.set_trajectory_fromarray(traj, cutoff_idx) for the batch (main/lib/mpc_with_speed.py:
        276-282): the speed reference drops to 0 from cutoff_idx[b] on; 999 (the reference's sentinel) or None = no cut."""
        if cutoff_idx is None:
            _lib.check(self._lib.mpcb_session_set_speed_cutoff_host(self._h, None), self._h, "mpcb_session_set_speed_cutoff_host")
            return
        c = np.ascontiguousarray(cutoff_idx, dtype=np.int32)
        assert c.shape == (self.B,)
        _lib.check(self._lib.mpcb_session_set_speed_cutoff_host(self._h, _p(c)), self._h, "mpcb_session_set_speed_cutoff_host")

    def set_scripted_obstacles(self, scripts):
        """scripts: per ego a list of (direction, turning, speed, offset) tuples -- the constructor arguments of the
        reference's `MovingObstacleTIntersection` (main/lib/moving_obstacles.py:165-196; offset None or <= 0 = starts
        at once) -- or a ready [B, MAX_OBSTACLES, 4] array with n_obs inferred from NaN rows."""
        B, M = self.B, _lib.MAX_OBSTACLES
        arr = np.zeros((B, M, 4))
        n = np.zeros(B, np.int32)
        if isinstance(scripts, np.ndarray):
            assert scripts.shape == (B, M, 4)
            n[:] = (~np.isnan(scripts[:, :, 0])).sum(axis=1)
            arr[:] = np.nan_to_num(scripts)
        else:
            assert len(scripts) == B
            for b, lst in enumerate(scripts):
                assert len(lst) <= M
                n[b] = len(lst)
                for o, (direction, turning, speed, offset) in enumerate(lst):
                    arr[b, o] = (direction, 1.0 if turning else 0.0, speed, 0.0 if offset is None else offset)
        _lib.check(self._lib.mpcb_session_set_scripted_host(self._h, _p(arr), _p(n)), self._h,
                   "mpcb_session_set_scripted_host")

    def scripted_loop_dev(self, nsteps: int, history_ptr: int = 0, obs_history_ptr: int = 0, cuda_stream: int = 0,
                          goal_test: bool = True):
        """The reference's whole loop body with scripted cross traffic (main/scenarios/mpc_intersection.py:95-159) on
        the device: nsteps x (is_goal | obstacle get()/step() | coupling | MPC.step | Simulation.step), two launches
        per control step, no host round trip.  obs_history_ptr: device float64 [nsteps,B,MAX_OBSTACLES,6] or 0."""
        _lib.check(self._lib.mpcb_session_scripted_loop_dev(
            self._h, int(nsteps), C.c_void_p(history_ptr) if history_ptr else None,
            C.c_void_p(obs_history_ptr) if obs_history_ptr else None, 0 if goal_test else 1, C.c_void_p(cuda_stream)),
            self._h, "mpcb_session_scripted_loop_dev")

    def coupling(self, state, route, agent_idx, ncourse, obs=None, n_obs=None, vehicles_per_scenario=0):
        """Row A9 with host buffers (mpcb_coupling_host): the coupling part of the reference loop body
        (main/scenarios/mpc_intersection.py:100-139) for B egos.  agent_idx / ncourse are updated in place.
        obs [B, MAX_OBSTACLES, 6] rows (x, y, v, yaw, a, steer), n_obs [B]."""
        B = len(state)
        for a, dt in ((state, np.float64), (route, np.int32), (agent_idx, np.int32), (ncourse, np.int32)):
            assert a.dtype == dt and a.flags.c_contiguous and len(a) == B
        if obs is not None:
            assert obs.dtype == np.float64 and obs.shape == (B, _lib.MAX_OBSTACLES, 6) and obs.flags.c_contiguous
            assert n_obs is not None and n_obs.dtype == np.int32 and n_obs.shape == (B,)
        out = dict(collided=np.zeros(B, np.int32), coll_idx=np.zeros(B, np.int32), coll_xy=np.zeros((B, 2)),
                   n_res=np.zeros(B, np.int32), flags=np.zeros(B, np.int32))
        _lib.check(self._lib.mpcb_coupling_host(
            self._h, B, _p(state), _p(route), _p(obs), _p(n_obs), _p(agent_idx), _p(ncourse), _p(out["collided"]),
            _p(out["coll_idx"]), _p(out["coll_xy"]), _p(out["n_res"]), _p(out["flags"]), int(vehicles_per_scenario)),
            self._h, "mpcb_coupling_host")
        return out

    # ---- attribute access -----------------------------------------------------------------------------------
    def _meta(self, name):
        fid, dt, tail = _FIELDS[name]
        return fid, dt, (self.B,) + tail(self.cfg.T)

    def fetch(self, name) -> np.ndarray:
        fid, dt, shape = self._meta(name)
        out = np.empty(shape, dt)
        _lib.check(self._lib.mpcb_session_fetch_host(self._h, fid, _p(out), out.nbytes), self._h, "mpcb_session_fetch_host")
        return out

    def store(self, name, value):
        fid, dt, shape = self._meta(name)
        v = np.ascontiguousarray(value, dtype=dt)
        assert v.shape == shape, (v.shape, shape)
        _lib.check(self._lib.mpcb_session_store_host(self._h, fid, _p(v), v.nbytes), self._h, "mpcb_session_store_host")

    def device_view(self, name) -> _DevView:
        fid, dt, shape = self._meta(name)
        ptr = self._lib.mpcb_session_device_ptr(self._h, fid)
        if not ptr:
            raise _lib.MpcbError(f"no device pointer for {name}")
        return _DevView(ptr, shape, dt)

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
