"""The C-ABI library loads on a CPU-only box and exports every symbol include/mpcb.h declares; without a
CUDA device the product fails loudly (no CPU fallback).  No compute calls here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import REPO


@pytest.fixture(scope="module")
def lib():
    from mpc_for_av_at_intersection_b200 import _lib
    _lib.build()
    return _lib.lib()


def _declared():
    src = open(os.path.join(REPO, "include", "mpcb.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mpcb_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported(lib):
    from mpc_for_av_at_intersection_b200 import _lib
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/mpcb.h but not exported by libmpcb.so"
    assert sorted(_lib.EXPORTS) == names, "the ctypes binding and the header must list the same entry points"


def test_version_and_default_config(lib):
    from mpc_for_av_at_intersection_b200 import _lib
    assert lib.mpcb_version() == 100
    c = _lib.mpcb_config()
    lib.mpcb_default_config(C.byref(c))
    # main/config/mpc_config.json + main/lib/simulation.py:23-25
    assert (c.T, c.max_iter) == (13, 1)
    assert (c.w_perp, c.w_para) == (20.0, 1.0)
    assert list(c.R) == [0.01, 0.01] and list(c.Rd) == [0.01, 1.0] and list(c.Q_v_yaw) == [0.0, 0.5]
    assert list(c.Qf) == [1.0, 1.0, 0.0, 0.5]
    assert c.max_steer == float(np.deg2rad(45.0)) and c.max_speed == 30.0 / 3.6 and c.min_speed == -5.0
    assert (c.max_accel, c.max_decel, c.max_dsteer_deg) == (2.0, -10.0, 30.0)
    assert c.min_ref_speed == 10 / 3.6 and c.dt == 0.2 and c.wheelbase == 2.86
    n = lib.mpcb_supported_horizons(None, 0)
    buf = (C.c_int * n)()
    lib.mpcb_supported_horizons(buf, n)
    assert {10, 13, 20, 40, 80} <= set(buf)


def test_smooth_yaw_matches_oracle(lib):
    import mpc_oracle as mo
    rng = np.random.default_rng(0)
    for _ in range(20):
        yaw = np.cumsum(rng.normal(0, 0.8, 200))
        yaw = (yaw + np.pi) % (2 * np.pi) - np.pi
        a, b = yaw.copy(), yaw.copy()
        lib.mpcb_smooth_yaw(a.ctypes.data_as(C.c_void_p), len(a))
        mo.smooth_yaw(b)
        assert np.array_equal(a, b)
        from mpc_for_av_at_intersection_b200.courses import smooth_yaw_host
        assert np.array_equal(smooth_yaw_host(yaw.copy()), a)       # the library-free twin used by bench.py's CPU arm


def test_no_cpu_fallback(lib):
    """Without a CUDA device mpcb_create must fail with MPCB_ENODEV and say why."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from mpc_for_av_at_intersection_b200 import CourseSet, MpcConfig, MpcSession, _lib, load_intersection_routes
    c = _lib.mpcb_config()
    lib.mpcb_default_config(C.byref(c))
    h = C.c_void_p()
    assert lib.mpcb_create(C.byref(c), 8, 0, C.byref(h)) == -3
    assert b"no CPU fallback" in lib.mpcb_last_error(None)
    with pytest.raises(_lib.MpcbError):
        MpcSession(MpcConfig(), CourseSet(load_intersection_routes()), 8)
    c.T = 17
    assert lib.mpcb_create(C.byref(c), 8, 0, C.byref(h)) == -1     # no kernel for this horizon: EINVAL
