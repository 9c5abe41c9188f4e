"""The oracle (numpy restatement, oracle/mpc_oracle.py) against the golden vectors recorded from the
UNMODIFIED reference (oracle/gen_golden.py): integer outputs bit-exact, controls to 1e-7.
A strided subsample keeps the CPU suite short (the dense float64 IPM costs ~0.1 s per solve)."""
import numpy as np
import pytest

import mpc_oracle as mo
from conftest import golden, load_routes


def _replay(g, k, routes_smooth, T):
    r = int(g["route"][k])
    nc = int(g["ncourse"][k])
    full = routes_smooth[r]
    L = float(g["L"][k]) if "L" in g.files else 2.86
    dl = float(np.linalg.norm(full[0, :2] - full[1, :2]))
    om = mo.OracleMPC(full[:nc, 0].copy(), full[:nc, 1].copy(), full[:nc, 2].copy(), dl, L, 0.2,
                      mo.Cfg.from_dict({"T": T}))
    om.goal = (full[-1, 0], full[-1, 1])
    om.target_ind = int(g["target_in"][k])
    if g["has_warm"][k]:
        om.oa, om.odelta = g["oa_in"][k].copy(), g["od_in"][k].copy()
    om.di, om.ai = float(g["di_in"][k]), float(g["ai_in"][k])
    raised = False
    try:
        om.step(g["state"][k])
    except Exception:
        raised = True
    return om, raised


@pytest.mark.parametrize("name,T,stride", [("random_T13.npz", 13, 3), ("random_T20.npz", 20, 4),
                                           ("closed_loop_T13.npz", 13, 5), ("closed_loop_T20.npz", 20, 6),
                                           ("random_T40.npz", 40, 8), ("closed_loop_T80.npz", 80, 10),
                                           ("edge_cases_T10.npz", 10, 1), ("edge_cases_T13.npz", 13, 2)])
def test_oracle_matches_reference_golden(name, T, stride, routes_smooth):
    g = golden(name)
    n = len(g["state"])
    worst = 0.0
    for k in range(0, n, stride):
        om, raised = _replay(g, k, routes_smooth, T)
        assert raised == bool(g["raised"][k])
        if raised:
            continue
        assert om.target_ind == int(g["target_out"][k])
        assert np.array_equal(om.xref, g["xref"][k]), "xref is gathered by integer indices: bit-exact"
        assert (om.odelta is not None) == bool(g["solved"][k])
        if g["solved"][k]:
            worst = max(worst, abs(om.di - g["di"][k]), abs(om.ai - g["ai"][k]))
            assert np.abs(om.oa - g["oa"][k]).max() < 1e-6 and np.abs(om.odelta - g["od"][k]).max() < 1e-6
            assert np.abs(om.ox - g["ox"][k]).max() < 1e-6 and np.abs(om.oy - g["oy"][k]).max() < 1e-6
        else:
            assert om.ai == -10.0 and om.di == g["di_in"][k]
    assert worst < 1e-7, worst


def test_condensed_model_agrees_with_oracle(routes_smooth):
    """The condensed formulation the CUDA kernel implements (oracle/condensed_model.py: closed-form
    sensitivities, suffix-moment Hessian, structured constraint rows, Mehrotra iteration) has the same optimum
    as the reference's sparse QP (strictly convex in u: unique), and its structured Hessian/gradient match
    dense finite differences of the cost."""
    import condensed_model as cm
    g = golden("random_T13.npz")
    T = 13
    checked = 0
    for k in range(0, len(g["state"]), 25):
        if not g["solved"][k]:
            continue
        om, raised = _replay(g, k, routes_smooth, T)
        assert not raised
        x, y, yaw, v = g["state"][k]
        c = cm.Condensed(om.xref, om.xbar, [x, y, v, yaw], om.reaches_end, om.cfg, om.L, 0.2)
        sol = c.solve()
        assert sol["status"] == 0
        # whole-horizon inputs: see the degeneracy note in tests/test_gpu_parity_step.py
        assert abs(sol["oa"][0] - om.oa[0]) < 1e-7 and abs(sol["od"][0] - om.odelta[0]) < 1e-7
        assert np.abs(sol["oa"] - om.oa).max() < 2e-4 and np.abs(sol["od"] - om.odelta).max() < 2e-4
        # Hessian blocks vs differences of the scan gradient (J is quadratic: exact up to rounding)
        A, B, C = c.hessian()
        H = np.block([[A, B.T], [B, C]])
        g0 = np.concatenate(c.grad(np.zeros(T), np.zeros(T)))
        for j in (0, 5, T - 1, T, T + 7, 2 * T - 1):
            e = np.zeros(2 * T)
            e[j] = 1.0
            gj = np.concatenate(c.grad(e[:T], e[T:]))
            assert np.abs((gj - g0) - H[:, j]).max() < 1e-8 * max(1.0, np.abs(H[:, j]).max())
        checked += 1
    assert checked >= 5


# ---- controller variants: oracle vs the reference's mpc_with_speed / mpc_sensitivity -------------
WITH_SPEED = {"T": 13, "w_perp": 10.0, "w_para": 1.0, "R": [0.01, 0.01], "Rd": [0.01, 1.0], "Q_v_yaw": [20.0, 0.5],
              "Qf": [1.0, 1.0, 0.0, 0.5], "GOAL_DIS": 1.5, "STOP_SPEED": 0.5 / 3.6, "MAX_DSTEER": 30.0, "MAX_ACCEL": 2.0,
              "MAX_DECEL": -5}          # main/lib/mpc_with_speed.py:16-36,161,165


def _replay_variant(g, k, routes_smooth, cfgd, ref_speed=None):
    r = int(g["route"][k])
    full = routes_smooth[r]
    om = mo.OracleMPC(full[:, 0].copy(), full[:, 1].copy(), full[:, 2].copy(),
                      float(np.linalg.norm(full[0, :2] - full[1, :2])), 2.86, 0.2, mo.Cfg.from_dict(cfgd))
    om.ref_speed = ref_speed
    om.set_trajectory_fromarray(full, int(g["cutoff_idx"][k]) if "cutoff_idx" in g.files else 999)
    om.target_ind = int(g["target_in"][k])
    if g["has_warm"][k]:
        om.oa, om.odelta = g["oa_in"][k].copy(), g["od_in"][k].copy()
    om.di, om.ai = float(g["di_in"][k]), float(g["ai_in"][k])
    om.step(g["state"][k])
    return om


def _check_variant(om, g, k):
    assert om.target_ind == int(g["target_out"][k])
    assert np.array_equal(om.xref, g["xref"][k])
    assert (om.odelta is not None) == bool(g["solved"][k])
    if g["solved"][k]:
        assert np.abs(om.oa - g["oa"][k]).max() < 1e-6 and np.abs(om.odelta - g["od"][k]).max() < 1e-6
        return max(abs(om.di - g["di"][k]), abs(om.ai - g["ai"][k]))
    return 0.0


def test_oracle_matches_reference_mpc_with_speed(routes_smooth):
    g = golden("variant_with_speed.npz")
    assert (g["cutoff_idx"] != 999).sum() > 50 and (g["xref"][:, 2] == 0).any() and (g["xref"][:, 2] > 6.9).any()
    worst = max(_check_variant(_replay_variant(g, k, routes_smooth, WITH_SPEED, 25 / 3.6), g, k)
                for k in range(0, len(g["state"]), 3))
    assert worst < 1e-7, worst


def test_oracle_matches_reference_mpc_sensitivity(routes_smooth):
    g = golden("variant_sensitivity.npz")
    assert int(g["n_cases"]) == 30
    worst = 0.0
    for k in range(0, len(g["state"]), 9):
        w = g["weights"][k]
        cfgd = {"T": 13, "w_perp": w[0], "w_para": w[1], "R": w[2:4], "Rd": w[4:6], "Q_v_yaw": w[6:8], "Qf": w[8:12]}
        worst = max(worst, _check_variant(_replay_variant(g, k, routes_smooth, cfgd), g, k))
    assert worst < 1e-7, worst


def test_condensed_jerk_model_matches_reference_golden():
    """oracle/condensed_jerk.py (the math of the CUDA jerk kernel: standard condensed model at the effective acceleration
    e = E [a; z0], pulled back through E) against steps recorded from the unmodified main/lib/mpc_jerk.py."""
    import condensed_jerk as cj
    g = golden("variant_jerk.npz")
    routes = load_routes(True)
    T = 13
    cfg = cj.jerk_cfg(T)
    worst, n = 0.0, 0
    for i in range(0, len(g["state"]), 9):
        if g["raised"][i] or not g["solved"][i]:
            continue
        r = int(g["route"][i])
        full = routes[r]
        dl = float(np.linalg.norm(full[0, :2] - full[1, :2]))
        full = full[: int(g["ncourse"][i])]
        x, y, yaw, v = g["state"][i]
        warm = bool(g["has_warm"][i])
        oa = np.nan_to_num(g["oa_in"][i]) if warm else np.zeros(T)
        od = np.nan_to_num(g["od_in"][i]) if warm else np.zeros(T)
        xref, tind, idx, re = mo.calc_ref_trajectory(x, y, v, full[:, 0], full[:, 1], full[:, 2], dl, 0.2, int(g["target_in"][i]), T, None, None)
        assert tind == int(g["target_out"][i])
        xbar = mo.predict_motion([x, y, v, yaw], list(oa), list(od), 2.86, 0.2, T)
        sol = cj.CondensedJerk(xref, xbar, [x, y, v, yaw], re, cfg, 2.86, 0.2).solve()
        worst = max(worst, np.abs(sol["oa"] - g["oa"][i]).max(), np.abs(sol["od"] - g["od"][i]).max(),
                    np.abs(sol["ov"] - g["ov"][i]).max(), np.abs(sol["ox"] - g["ox"][i]).max())
        n += 1
    assert n > 30 and worst < 1e-8, (n, worst)
