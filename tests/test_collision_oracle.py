"""Row A9 oracle (oracle/collision_oracle.py) against vectors recorded from the reference's own functions:
tests/golden/collision_random.npz (oracle/gen_golden_collision.py) and the scripted-obstacle closed loop inside
tests/golden/closed_loop_T{13,20}.npz.  Integer outputs and flags bit-exact; the collision point is a path point,
hence exact too."""
import numpy as np
import pytest

import collision_oracle as co
from conftest import golden

CARS = [(2.86, 2.0, 3.5), (4.0, 2.04, 4.84)]       # wheelbase, width, length (main/lib/car_dimensions.py:82-107)


def _check(out, g, k):
    assert out["n_res"] == int(g["n_res"][k])
    assert out["collided"] == int(g["collided"][k])
    assert out["cutoff_idx"] == int(g["cutoff_idx"][k])
    if out["collided"]:
        assert out["coll_idx"] == int(g["coll_idx"][k])
        assert out["coll_x"] == g["coll_x"][k] and out["coll_y"] == g["coll_y"][k]


def test_random_scenes(routes_smooth):
    g = golden("collision_random.npz")
    n = len(g["route"])
    assert 0.2 < g["collided"].mean() < 0.9
    for k in range(n):
        L, w, ln = CARS[int(g["car"][k])]
        offs, radius = co.circle_offsets(L, w, ln)
        obs = g["obs"][k][:int(g["n_obs"][k])]
        out = co.coupling_step(routes_smooth[int(g["route"][k])], int(g["agent_idx"][k]), float(g["v"][k]), obs, offs,
                               radius, L)
        _check(out, g, k)


@pytest.mark.parametrize("T", [13, 20])
def test_scripted_obstacle_closed_loop(T, routes_smooth):
    g = golden(f"closed_loop_T{T}.npz")
    sel = np.flatnonzero(g["loop"] == 12)
    assert g["collided"][sel].sum() > 10
    L, w, ln = CARS[0]
    offs, radius = co.circle_offsets(L, w, ln)
    full = routes_smooth[int(g["route"][sel[0]])]
    for k in sel:
        out = co.coupling_step(full, int(g["agent_idx"][k]), float(g["state"][k][3]), g["obs"][k], offs, radius, L)
        _check(out, g, k)
        assert int(g["ncourse"][k]) == out["cutoff_idx"]         # what set_trajectory_fromarray received


def test_no_obstacles_and_degenerate_paths(routes_smooth):
    offs, radius = co.circle_offsets(*CARS[0])
    full = routes_smooth[0]
    out = co.coupling_step(full, 10, 3.0, np.zeros((0, 6)), offs, radius, 2.86)
    assert out["collided"] == 0 and out["cutoff_idx"] == len(full)
    last = co.coupling_step(full, len(full) - 1, 0.0, np.array([[full[-1, 0], full[-1, 1], 0, 0, 0, 0.0]]), offs, radius, 2.86)
    assert last["n_res"] == 1 and last["collided"] == 1 and last["cutoff_idx"] == len(full)   # max(idx + 1, j - 72)


REF_SCRIPTS = [(1, False, 25 / 3.6, 2.), (-1, True, 25 / 3.6, 4.)]        # main/scenarios/mpc_intersection.py:41-44


@pytest.mark.parametrize("T", [13, 20])
def test_scripted_obstacle_model_is_bit_exact(T):
    """oracle ScriptedObstacle vs every `o.get()` row the reference's MovingObstacleTIntersection objects returned."""
    g = golden(f"closed_loop_T{T}.npz")
    sel = np.flatnonzero(g["loop"] == 12)
    obs = [co.ScriptedObstacle(*s, 2.86, 0.2) for s in REF_SCRIPTS]
    for k in sel:
        assert np.array_equal(np.array([o.get() for o in obs], float), g["obs"][k])
        for o in obs:
            o.step()
    assert g["obs"][sel][:, 1, 5].max() == 0.19 and g["obs"][sel][0, :, 2].max() == 0.0      # turning seen, both start parked


def test_scripted_closed_loop_oracle_reproduces_reference_loop(routes_smooth):
    """The composed oracle loop (ScriptedObstacle + coupling_step + OracleMPC + plant) against the reference's own
    recorded loop with two scripted cars: steps to goal, every cut length and obstacle row bit-exact; states to 1e-6
    over the first 45 steps (the route is a left turn: see DESIGN.md "closed-loop sensitivity")."""
    T = 13
    g = golden(f"closed_loop_T{T}.npz")
    sel = np.flatnonzero(g["loop"] == 12)
    offs, radius = co.circle_offsets(*CARS[0])
    full = routes_smooth[int(g["route"][sel[0]])]
    hist, nclog, obslog, n = co.scripted_closed_loop(full, g["state"][sel[0]], REF_SCRIPTS, len(sel) + 3, T, offs, radius, 2.86)
    assert n == len(sel)
    assert np.array_equal(nclog[:n], g["ncourse"][sel])
    assert np.array_equal(obslog[:n], g["obs"][sel])
    off = g["history_off"]
    ref = g["history"][off[12]:off[13]][1:n + 1]
    assert np.abs(hist[:45, :4] - ref[:45, :4]).max() < 1e-6
    assert np.abs(hist[:n, :2] - ref[:, :2]).max() < 1e-3
