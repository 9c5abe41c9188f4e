"""SURVEY 8f N2 on the GPU: the batched A* planner (mpcb_plan_routes_host through the C ABI, one warp per search) against
the 84 searches recorded from the unmodified reference planner and against the CPU oracle on fresh scenarios.
Integer outputs (primitive index sequence, path length) and the cost are bit-exact class."""
import dataclasses

import numpy as np
import pytest

import astar_oracle as ao
from conftest import golden, load_routes

pytestmark = pytest.mark.gpu


class _Car:
    def __init__(self, g):
        self.circle_centers = g["circle_centers"]
        self.radius = float(g["radius"])


def _setup():
    from mpc_for_av_at_intersection_b200 import planner
    g = golden("astar_intersection.npz")
    lib = planner.PrimitiveLibrary(planner.load_motion_primitives("bicycle_model"), _Car(g))
    return planner, g, lib


def test_all_recorded_searches_in_one_launch():
    planner, g, lib = _setup()
    n = len(g["cost"])
    base = {}
    scs = []
    for k in range(n):
        key = (int(g["start_pos"][k]), int(g["turn"][k]))
        if key not in base:
            base[key] = planner.intersection(turn_indicator=key[1], start_pos=key[0])
        scs.append(dataclasses.replace(base[key], start=tuple(g["start"][k])))
    r = planner.plan_batch(scs, lib, margin=float(g["radius"]), max_path=24)
    same_count = 0
    for k in range(n):
        if g["failed"][k]:
            assert r.status[k] == planner.PLAN_NO_SOLUTION and r.n_path[k] == 0 and np.isnan(r.cost[k])
            continue
        assert r.status[k] == planner.PLAN_OK, (k, r.status[k])
        m = int(g["n_path"][k])
        assert r.n_path[k] == m and np.array_equal(r.mp_seq[k], g["mp_seq"][k]), (k, r.mp_seq[k], g["mp_seq"][k])
        assert r.cost[k] == g["cost"][k]
        assert np.abs(r.nodes[k, :m] - g["nodes"][k][:m]).max() < 1e-9
        assert r.n_traj[k] == g["n_traj"][k] and np.abs(r.trajectory(k) - g["traj"][k][:r.n_traj[k]]).max() < 1e-9
        assert abs(int(r.n_expanded[k]) - int(g["n_expanded"][k])) <= 0.25 * int(g["n_expanded"][k]) + 2
        same_count += int(r.n_expanded[k] == g["n_expanded"][k])
    print(f"{n} searches (max {g['n_expanded'].max()} expansions): sequences, path lengths and costs identical; "
          f"expansion counts identical in {same_count}")


def test_fresh_scenarios_match_oracle():
    """Random start poses on random routes (not in the golden file) against the fixed-order CPU oracle: the same IEEE
    operations in the same order on both sides (including sin / cos), so sequences, costs, nodes AND expansion counts are
    identical -- every tie of the search falls the same way."""
    planner, g, lib = _setup()
    rng = np.random.default_rng(99)
    scs = []
    for _ in range(48):
        sc = planner.intersection(turn_indicator=int(rng.integers(1, 4)), start_pos=int(rng.integers(1, 5)))
        x, y, th = sc.start
        lon, lat, dth = rng.uniform(-1.0, 4.0), rng.normal(0, 0.2), rng.normal(0, 0.02)
        scs.append(dataclasses.replace(sc, start=(x + np.cos(th) * lon - np.sin(th) * lat, y + np.sin(th) * lon + np.cos(th) * lat, th + dth)))
    r = planner.plan_batch(scs, lib, margin=float(g["radius"]))
    exact = checked = 0
    for k, sc in enumerate(scs):
        if r.n_expanded[k] > 1500:
            continue                                     # keep the CPU side of the test short
        hp = [o.to_convex(margin=float(g["radius"])) for o in sc.obstacles]
        o = ao.AStarOracle(sc.start, sc.goal_point, [*sc.goal_area.xy1, *sc.goal_area.xy2], sc.allowed_goal_theta_difference, hp,
                           lib.points, lib.n_points, lib.length, lib.cc)
        res = o.run()
        checked += 1
        if res is None:
            assert r.status[k] == planner.PLAN_NO_SOLUTION
            continue
        cost, path, seq = res
        assert r.status[k] == planner.PLAN_OK and list(r.mp_seq[k, :r.n_path[k] - 1]) == seq and r.cost[k] == cost, k
        assert np.array_equal(r.nodes[k, :len(path)], np.array(path)), "node coordinates must be bit-identical"
        exact += int(r.n_expanded[k] == o.n_expanded)
    assert checked >= 30 and exact == checked, (checked, exact)
    print(f"{checked} fresh searches: sequences and costs identical, expansion counts identical in {exact}")


def test_planner_facade_feeds_the_mpc():
    """`MotionPrimitiveSearch(scenario, car, mps, margin).run()` as main/scenarios/mpc_intersection.py:56-60 calls it: the
    trajectory equals the route the reference planner produced (shipped route data), and drives the MPC facade."""
    planner, g, lib = _setup()
    car = _Car(g)
    mps = planner.load_motion_primitives(version="bicycle_model")
    search = planner.MotionPrimitiveSearch(planner.intersection(start_pos=4, turn_indicator=1), car, mps, margin=car.radius)
    cost, path, trajectory_full = search.run(debug=False)
    ref = load_routes(False)[9]                           # route (4, 1)
    assert trajectory_full.shape == ref.shape and np.abs(trajectory_full - ref).max() < 1e-9
    assert abs(cost - 54.78) < 1e-9 and search.primitive_names[:4] == ["straight", "straight", "straight", "left1"]
    with pytest.raises(Exception, match="No solution found"):
        bad = dataclasses.replace(planner.intersection(start_pos=3, turn_indicator=1), start=tuple(g["start"][np.flatnonzero(g["failed"])[0]]))
        planner.MotionPrimitiveSearch(bad, car, mps, margin=car.radius).run()


def test_capacity_overflow_is_reported_not_silent():
    """A search that needs more open-list or closed-set entries than it was given must come back as PLAN_OVERFLOW (with no
    path), never as a wrong path; the same search with room succeeds."""
    planner, g, lib = _setup()
    k = int(np.argmax(g["n_expanded"] * (g["failed"] == 0)))          # the longest recorded search
    sc = dataclasses.replace(planner.intersection(turn_indicator=int(g["turn"][k]), start_pos=int(g["start_pos"][k])),
                             start=tuple(g["start"][k]))
    small = planner.plan_batch([sc], lib, margin=float(g["radius"]), max_open=512, max_closed=256, grow=0)
    assert small.status[0] == planner.PLAN_OVERFLOW and small.n_path[0] == 0 and (small.mp_seq[0] == -1).all() and np.isnan(small.cost[0])
    ok = planner.plan_batch([sc], lib, margin=float(g["radius"]))
    assert ok.status[0] == planner.PLAN_OK and ok.cost[0] == g["cost"][k]
    # the capacities are a first guess: with growth on, the same tiny start is re-run at 4x, 16x, ... and ends with the same answer
    grown = planner.plan_batch([sc], lib, margin=float(g["radius"]), max_open=512, max_closed=256, grow=4)
    assert grown.status[0] == planner.PLAN_OK and grown.cost[0] == g["cost"][k]
    assert np.array_equal(grown.mp_seq[0], ok.mp_seq[0]) and grown.n_expanded[0] == ok.n_expanded[0]
