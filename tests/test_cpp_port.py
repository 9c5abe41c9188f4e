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


@pytest.mark.parametrize("name,T", [("random_T13.npz", 13), ("random_T20.npz", 20), ("closed_loop_T20.npz", 20), ("random_T40.npz", 40),
                                    ("edge_cases_T10.npz", 10), ("edge_cases_T13.npz", 13), ("edge_cases_T20.npz", 20)])     # edge cases: raised / infeasible / cut paths
def test_cpp_port_matches_reference_golden(name, T):
    import mpc_port
    from mpc_for_av_at_intersection_b200 import MpcConfig
    g = golden(name)
    cs = _courses()
    cfgvec = mpc_port.pack_config(MpcConfig.from_json(T=T))
    n = len(g["state"])
    sel = np.arange(0, n, 1 if name.startswith("edge") else (2 if T <= 20 else 3))
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


def test_hard_cases_are_reproducible_and_solved_by_the_twin():
    """tests/golden/hard_cases.npz (oracle/gen_golden_hard_cases.py): the port's recorded answers are reproduced here, and
    the numpy twin of the kernel's interior point (oracle/condensed_model.py, gradient-scaled centred start) solves the
    T = 20 / 40 cases and two of the T = 80 ones -- in a third of the iterations the lam = 1 start needed."""
    import mpc_port
    import condensed_model as cm
    import mpc_oracle as mo
    from mpc_for_av_at_intersection_b200 import MpcConfig
    g = golden("hard_cases.npz")
    cs = _courses()
    routes = load_routes(True)
    for T in (20, 40, 80):
        sel = np.flatnonzero(g["T"] == T)[:2]
        c = lambda k, dt: np.ascontiguousarray(g[k][sel]).astype(dt)
        tgt, warm, di_ai = c("target_ind", np.int32), c("warm", np.int32), c("di_ai", np.float64)
        oa, od = np.ascontiguousarray(g["oa"][sel][:, :T]), np.ascontiguousarray(g["odelta"][sel][:, :T])
        oa0, od0 = oa.copy(), od.copy()
        out = mpc_port.step_batch(mpc_port.pack_config(MpcConfig.from_json(T=T)), cs.packed, cs.dl, c("state", np.float64), c("route", np.int32),
                                  c("ncourse", np.int32), np.full(len(sel), 2.86), tgt, warm, oa, od, di_ai, want_traj=False)
        assert np.array_equal(out["status"], g["port_status"][sel]) and np.abs(di_ai - g["port_di_ai"][sel]).max() < 1e-9
        cfg = mo.Cfg.from_dict({"T": T})
        for j, i in enumerate(sel):
            full = routes[int(g["route"][i])]
            dl = float(np.linalg.norm(full[0, :2] - full[1, :2]))
            full = full[: int(g["ncourse"][i])]
            x, y, yaw, v = g["state"][i]
            xref, tind, idx, re = mo.calc_ref_trajectory(x, y, v, full[:, 0], full[:, 1], full[:, 2], dl, 0.2, int(g["target_ind"][i]), T, None, None)
            xbar = mo.predict_motion([x, y, v, yaw], list(oa0[j]), list(od0[j]), 2.86, 0.2, T)
            r = cm.Condensed(xref, xbar, [x, y, v, yaw], re, cfg, 2.86, 0.2).solve(oa0[j], od0[j], max_iter=100)
            assert r["status"] == 0 and r["iters"] < 0.7 * g["port_iters"][i] + 12, (T, i, r["iters"])
            assert max(abs(r["od"][0] - g["port_di_ai"][i, 0]), abs(r["oa"][0] - g["port_di_ai"][i, 1])) < 1e-4
