"""The plain C++ CPU restatement (oracle/mpc_port.cpp: generic dense condensing + dense Mehrotra IPM) against the golden
vectors recorded from the unmodified reference -- a second CPU path, algorithmically independent of both the numpy
oracle (sparse form, null-space IPM + active-set polish) and the CUDA kernels (closed-form condensing, structured
operators).  Integer outputs bit-exact, applied controls 1e-6 (its IPM stops at the kernel's tolerances, no polish)."""
import os
import sys

import numpy as np
import pytest

from conftest import golden, load_routes

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))


def _courses():
    from mpc_for_av_at_intersection_b200.courses import CourseSet
    return CourseSet(load_routes(False))


@pytest.mark.parametrize("name,T", [("random_T13.npz", 13), ("random_T20.npz", 20), ("closed_loop_T20.npz", 20), ("random_T40.npz", 40)])
def test_cpp_port_matches_reference_golden(name, T):
    import mpc_port
    from mpc_for_av_at_intersection_b200 import MpcConfig
    g = golden(name)
    cs = _courses()
    cfgvec = mpc_port.pack_config(MpcConfig.from_json(T=T))
    n = len(g["state"])
    sel = np.arange(0, n, 2 if T <= 20 else 3)
    state = np.ascontiguousarray(g["state"][sel])
    route = g["route"][sel].astype(np.int32)
    ncourse = g["ncourse"][sel].astype(np.int32)
    target = g["target_in"][sel].astype(np.int32)
    warm = g["has_warm"][sel].astype(np.int32)
    oa = np.nan_to_num(g["oa_in"][sel]).copy()
    od = np.nan_to_num(g["od_in"][sel]).copy()
    di_ai = np.stack([g["di_in"][sel], g["ai_in"][sel]], axis=1).copy()
    L = np.ascontiguousarray(g["L"][sel]) if "L" in g.files else np.full(len(sel), 2.86)
    out = mpc_port.step_batch(cfgvec, cs.packed, cs.dl, state, route, ncourse, L, target, warm, oa, od, di_ai)
    raised = g["raised"][sel].astype(bool)
    assert np.array_equal(out["status"] == 3, raised)
    ok = ~raised
    assert np.array_equal(target[ok], g["target_out"][sel][ok])
    assert np.array_equal(out["xref"][ok], g["xref"][sel][ok]), "xref is gathered by integer indices: bit-exact"
    solved = g["solved"][sel].astype(bool)
    assert np.array_equal((out["status"] <= 1) & ok, solved)
    s = solved
    e0 = max(np.abs(di_ai[s, 0] - g["di"][sel][s]).max(), np.abs(di_ai[s, 1] - g["ai"][sel][s]).max())
    eu = max(np.abs(oa[s] - g["oa"][sel][s]).max(), np.abs(od[s] - g["od"][sel][s]).max())
    ex = max(np.abs(out["traj"][s, 0] - g["ox"][sel][s]).max(), np.abs(out["traj"][s, 1] - g["oy"][sel][s]).max())
    print(f"{name}: n={len(sel)} max|du0|={e0:.2e} max|du|={eu:.2e} max|dxy|={ex:.2e} iters mean={out['iters'].mean():.1f}")
    assert e0 < 1e-6 and eu < 2e-4 and ex < 1e-5
    f = ~solved & ok
    assert np.all(di_ai[f, 1] == -10.0) and np.array_equal(di_ai[f, 0], g["di_in"][sel][f])
