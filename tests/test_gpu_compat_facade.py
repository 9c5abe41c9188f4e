"""The reference-shaped facades (compat.mpc / compat.mpc_with_speed / compat.mpc_sensitivity: one `MPC` object = a B=1
view of the batched engine) driven the way the reference's scenario scripts drive them, against the loops the
unmodified reference recorded.  Tolerance on the applied controls: 1e-6 (north_star: 1e-4)."""
import json
import os

import numpy as np
import pytest

from conftest import golden, load_routes

pytestmark = pytest.mark.gpu


def _car():
    from mpc_for_av_at_intersection_b200.compat.car_dimensions import BicycleModelDimensions
    return BicycleModelDimensions(skip_back_circle_collision_checking=False)


def test_mpc_facade_closed_loop_like_mpc_intersection():
    """main/scenarios/mpc_intersection.py:69-159 without obstacles, through compat.mpc + compat.simulation."""
    from mpc_for_av_at_intersection_b200.compat import mpc as cm
    from mpc_for_av_at_intersection_b200.compat.simulation import HistorySimulation, State
    if cm.T != 13:
        pytest.skip("compat.mpc picked up a non-default config")
    g = golden("closed_loop_T13.npz")
    loop = 3
    rid = int(g["loop_route"][loop])
    full = load_routes(False)[rid]
    dl = np.linalg.norm(full[0, :2] - full[1, :2])
    mpc = cm.MPC(cx=full[:, 0], cy=full[:, 1], cyaw=full[:, 2], dl=dl, dt=0.2, car_dimensions=_car())
    state = State(x=full[0, 0], y=full[0, 1], yaw=full[0, 2], v=0.0)
    sim = HistorySimulation(car_dimensions=_car(), sample_time=0.2, initial_state=state)
    n = 0
    while not mpc.is_goal(state) and n < 200:
        mpc.set_trajectory_fromarray(full)
        delta, acc = mpc.step(state)
        state = sim.step(a=acc, delta=delta, xref_deviation=mpc.get_current_xref_deviation())
        n += 1
    assert n == int(g["loop_len"][loop])
    off = g["history_off"]
    ref = g["history"][off[loop]:off[loop + 1]]
    h = sim.history
    got = np.array([h.x, h.y, h.yaw, h.v, h.delta, h.a, h.xref_deviation]).T
    assert got.shape == ref.shape
    assert np.abs(got[:, :4] - ref[:, :4]).max() < 1e-6 and np.abs(got[:, 4:6] - ref[:, 4:6]).max() < 1e-6


def test_mpc_with_speed_facade_replays_new_ref_scenario():
    """compat.mpc_with_speed.MPC driven like main/scenarios/mpc_intersection_new_ref.py:65-150: constructor with a
    30/3.6 cv, set_trajectory_fromarray(full, cutoff_idx) before every step, warm start carried by the object."""
    from mpc_for_av_at_intersection_b200.compat import mpc_with_speed as cm
    from mpc_for_av_at_intersection_b200.compat.simulation import State
    g = golden("variant_with_speed.npz")
    sel = np.flatnonzero(g["kind"] == 0)
    full = load_routes(False)[int(g["route"][sel[0]])]
    dl = np.linalg.norm(full[0, :2] - full[1, :2])
    mpc = cm.MPC(cx=full[:, 0], cy=full[:, 1], cv=np.full(len(full), 30 / 3.6), cyaw=full[:, 2], dl=dl, dt=0.2,
                 car_dimensions=_car())
    worst = 0.0
    for k in sel:
        x, y, yaw, v = g["state"][k]
        mpc.set_trajectory_fromarray(full, cutoff_idx=int(g["cutoff_idx"][k]))
        delta, acc = mpc.step(State(x=x, y=y, yaw=yaw, v=v))
        assert mpc.target_ind == int(g["target_out"][k])
        assert np.array_equal(mpc.xref, g["xref"][k])
        worst = max(worst, abs(delta - g["di"][k]), abs(acc - g["ai"][k]))
    assert cm.MAX_DECEL == -5 and mpc.cv[0] == cm.MAX_SPEED
    print(f"mpc_with_speed facade: {len(sel)} steps, max|du0| = {worst:.2e}")
    assert worst < 1e-6


def test_mpc_sensitivity_facade_rereads_its_json(tmp_path, monkeypatch):
    """compat.mpc_sensitivity.MPC re-reads the sensitivity JSON on every solve, as the reference does, so the sweep
    script's pattern (rewrite the file, build a new MPC, run) works unchanged."""
    cfg_file = tmp_path / "mpc_config_sensitivity.json"
    base = json.load(open(os.path.join(os.path.dirname(__file__), "..", "mpc_for_av_at_intersection_b200", "config",
                                       "mpc_config.json")))
    cfg_file.write_text(json.dumps(base))
    monkeypatch.setenv("MPCB_CONFIG_SENSITIVITY", str(cfg_file))
    from mpc_for_av_at_intersection_b200.compat import mpc_sensitivity as cm
    from mpc_for_av_at_intersection_b200.compat.simulation import State
    g = golden("variant_sensitivity.npz")
    worst = 0.0
    for case in (0, 14, 29):
        sel = np.flatnonzero(g["loop"] == case)
        w = g["weights"][sel[0]]
        cfg = dict(base, w_perp=w[0], w_para=w[1], R=list(w[2:4]), Rd=list(w[4:6]), Q_v_yaw=list(w[6:8]), Qf=list(w[8:12]))
        cfg_file.write_text(json.dumps(cfg))
        full = load_routes(False)[int(g["route"][sel[0]])]
        mpc = cm.MPC(cx=full[:, 0], cy=full[:, 1], cyaw=full[:, 2], dl=np.linalg.norm(full[0, :2] - full[1, :2]), dt=0.2,
                     car_dimensions=_car())
        for k in sel:
            x, y, yaw, v = g["state"][k]
            mpc.set_trajectory_fromarray(full)
            delta, acc = mpc.step(State(x=x, y=y, yaw=yaw, v=v))
            worst = max(worst, abs(delta - g["di"][k]), abs(acc - g["ai"][k]))
    print(f"mpc_sensitivity facade: 3 settings x 25 steps, max|du0| = {worst:.2e}")
    assert worst < 1e-6


def test_mpc_jerk_facade_closed_loop():
    """compat.mpc_jerk.MPC driven like the loop body of main/scenarios/mpc_intersection.py:95-159 (no obstacles) against the
    loop the unmodified main/lib/mpc_jerk.py recorded: steps to goal identical, controls 1e-6 over the first 40 steps,
    positions 1e-3 m to the goal."""
    from mpc_for_av_at_intersection_b200.compat import mpc_jerk as cm
    from mpc_for_av_at_intersection_b200.compat.simulation import HistorySimulation, State
    g = golden("variant_jerk.npz")
    sel = np.flatnonzero((g["kind"] == 0) & (g["loop"] == 0))
    full = load_routes(False)[int(g["route"][sel[0]])]
    dl = np.linalg.norm(full[0, :2] - full[1, :2])
    mpc = cm.MPC(cx=full[:, 0], cy=full[:, 1], cyaw=full[:, 2], dl=dl, dt=0.2, car_dimensions=_car())
    state = State(x=full[0, 0], y=full[0, 1], yaw=full[0, 2], v=0.0)
    sim = HistorySimulation(car_dimensions=_car(), sample_time=0.2, initial_state=state)
    n = 0
    while not mpc.is_goal(state) and n < 400:
        delta, acc = mpc.step(state)
        state = sim.step(a=acc, delta=delta, xref_deviation=mpc.get_current_xref_deviation())
        n += 1
    assert n == int(g["loop_len"][0]) == len(sel)
    h = sim.history
    got = np.array([h.x, h.y, h.yaw, h.v, h.delta, h.a]).T
    ref = g["history"][: n + 1]
    assert got.shape[0] == ref.shape[0]
    print(f"mpc_jerk facade: {n} steps, max|dpos|={np.abs(got[:, :2] - ref[:, :2]).max():.2e} max|du| first 40={np.abs(got[:41, 4:6] - ref[:41, 4:6]).max():.2e}")
    assert np.abs(got[:, :2] - ref[:, :2]).max() < 1e-3 and np.abs(got[:41, 4:6] - ref[:41, 4:6]).max() < 1e-6
    assert cm.NX == 5 and cm.MAX_DECEL == -5
