"""TEST INFRASTRUCTURE -- loader for oracle/mpc_port.cpp (the plain C++ CPU restatement of the MPC step).

Only tests/, __graft_entry__.smoke()/build() and bench.py's CPU legs may import this module."""
import ctypes as C
import hashlib
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "mpc_port.cpp")
_lib = None


def cpu_model():
    """(model name, flags hash) of the host CPU: the port is built -march=native, so a library built on another CPU
    model must not be reused (the build container and the GPU box differ)."""
    name, flags = "unknown", ""
    try:
        for ln in open("/proc/cpuinfo"):
            if ln.startswith("model name") and name == "unknown":
                name = ln.split(":", 1)[1].strip()
            elif ln.startswith("flags") and not flags:
                flags = ln.split(":", 1)[1].strip()
            if name != "unknown" and flags:
                break
    except OSError:
        pass
    return name, hashlib.sha1((name + flags).encode()).hexdigest()[:10]


LIB = os.path.join(HERE, "_build", f"libmpcport.{cpu_model()[1]}.so")


def build(force=False):
    """g++ -O3 -march=native -shared (no external libraries), one library per host CPU model; returns its path."""
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(SRC):
        os.makedirs(os.path.dirname(LIB), exist_ok=True)
        tmp = LIB + f".{os.getpid()}.tmp"
        cmd = ["g++", "-O3", "-march=native", "-shared", "-fPIC", "-std=c++17", "-pthread", SRC, "-o", tmp]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError("g++ failed:\n" + res.stderr[-3000:])
        os.replace(tmp, LIB)
    return LIB


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.mpcport_step_batch.restype = C.c_int
    return _lib


def pack_config(cfg):
    """cfg: mpc_for_av_at_intersection_b200.MpcConfig (host-side dataclass; no CUDA involved)."""
    T = int(cfg.T)
    return np.array([T, cfg.MAX_ITER, cfg.w_perp, cfg.w_para, cfg.R[0], cfg.R[1], cfg.Rd[0], cfg.Rd[1], cfg.Q_v_yaw[0],
                     cfg.Q_v_yaw[1], cfg.Qf[0] * T, cfg.Qf[1] * T, cfg.Qf[2] * T, cfg.Qf[3] * T, np.deg2rad(cfg.MAX_DSTEER),
                     cfg.MAX_ACCEL, cfg.MAX_DECEL, np.deg2rad(45.0), 30.0 / 3.6, -5.0, 10.0, 10.0 / 3.6, 0.2, 0.0], np.float64)


def step_batch(cfgvec, courses, dl, state, route, ncourse, L, target_ind, warm, oa, od, di_ai, want_traj=True, nthreads=0):
    """One MPC.step for B problems on the host cores.  Arrays as in the C ABI of the product (float64 / int32,
    C-contiguous); target_ind, warm, oa, od, di_ai are updated in place.  Returns dict(status, iters, traj, xref)."""
    B = len(state)
    T = int(cfgvec[0])
    ncmax = courses.shape[2]
    for a, dt in ((courses, np.float64), (dl, np.float64), (state, np.float64), (route, np.int32), (ncourse, np.int32), (L, np.float64),
                  (target_ind, np.int32), (warm, np.int32), (oa, np.float64), (od, np.float64), (di_ai, np.float64)):
        assert a.dtype == dt and a.flags.c_contiguous, (a.dtype, dt)
    status = np.zeros(B, np.int32)
    iters = np.zeros(B, np.int32)
    traj = np.zeros((B, 4, T + 1)) if want_traj else None
    xref = np.zeros((B, 4, T + 1)) if want_traj else None
    p = lambda a: a.ctypes.data_as(C.c_void_p) if a is not None else None
    lib().mpcport_step_batch(p(cfgvec), p(courses), p(dl), C.c_int(ncmax), C.c_int(B), p(state), p(route), p(ncourse), p(L),
                             p(target_ind), p(warm), p(oa), p(od), p(traj), p(xref), p(di_ai), p(status), p(iters), C.c_int(nthreads))
    return dict(status=status, iters=iters, traj=traj, xref=xref)
