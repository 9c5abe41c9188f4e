"""TEST INFRASTRUCTURE -- golden fixtures for the motion-primitive A* planner, recorded by running
the UNMODIFIED reference planner (main/lib/motion_primitive_search_modified.py + main/lib/a_star.py on
main/envs/intersection.py) through oracle/refharness.py.  Build container only:

    python oracle/gen_golden_astar.py

Writes
  mpc_for_av_at_intersection_b200/data/motion_primitives.npz
      the reference's primitive library (main/data/motion_primitives_{bicycle_model,prius}/*.pkl) as plain arrays:
      names (sorted file order = the dict order the search iterates in), points [9,61,3], total_length [9]
  tests/golden/astar_intersection.npz
      per case the scenario as numbers (start, goal point, goal box, allowed heading error, obstacle half-planes with
      the planner's margin) and what the reference returned: cost, primitive index sequence, path nodes, number of
      expanded nodes, full trajectory.  Cases: the 12 (start_pos, turn) routes and randomised start poses.
"""
import dataclasses
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
sys.path.insert(0, HERE)
import refharness as rh  # noqa: E402

MAX_OBS, MAX_HP, MAX_PATH = 32, 8, 24


def main(n_random=72, seed=2024):
    rh.setup({})
    from envs.intersection import intersection
    from lib.car_dimensions import BicycleModelDimensions
    from lib.motion_primitive import load_motion_primitives
    from lib.motion_primitive_search_modified import MotionPrimitiveSearch
    lib = {}
    for version in ("bicycle_model", "prius"):
        mps = load_motion_primitives(version=version)
        names = list(mps)
        lib[version + "_names"] = np.array(names)
        npts = np.array([len(mps[n].points) for n in names], np.int32)
        pts = np.full((len(names), int(npts.max()), 3), np.nan)
        for i, n in enumerate(names):
            pts[i, :npts[i]] = mps[n].points
        lib[version + "_points"] = pts                      # NaN-padded to the longest primitive
        lib[version + "_n_points"] = npts
        lib[version + "_total_length"] = np.array([float(mps[n].total_length) for n in names])
    np.savez_compressed(os.path.join(REPO, "mpc_for_av_at_intersection_b200", "data", "motion_primitives.npz"), **lib)

    car = BicycleModelDimensions(skip_back_circle_collision_checking=False)
    mps = load_motion_primitives(version="bicycle_model")
    names = list(mps)
    rng = np.random.default_rng(seed)
    cases = [(s, t, None) for s in (1, 2, 3, 4) for t in (1, 2, 3)]
    for _ in range(n_random):
        cases.append((int(rng.integers(1, 5)), int(rng.integers(1, 4)),
                      (rng.uniform(-2.0, 3.0), rng.normal(0, 0.25), rng.normal(0, 0.03))))
    rows = {k: [] for k in ("start", "goal", "goal_box", "allowed", "hp", "n_hp", "n_obs", "cost", "n_path", "mp_seq", "nodes",
                            "n_expanded", "traj", "n_traj", "start_pos", "turn", "failed")}
    for ci, (s, t, pert) in enumerate(cases):
        sc = intersection(start_pos=s, turn_indicator=t)
        if pert is not None:
            x, y, th = sc.start
            lon, lat, dth = pert
            sc = dataclasses.replace(sc, start=(x + np.cos(th) * lon - np.sin(th) * lat, y + np.sin(th) * lon + np.cos(th) * lat,
                                                th + dth))
        search = MotionPrimitiveSearch(sc, car, mps, margin=car.radius)
        hp = np.zeros((MAX_OBS, MAX_HP, 3))
        n_hp = np.zeros(MAX_OBS, np.int32)
        assert len(search._obstacles_hp) <= MAX_OBS
        for i, h in enumerate(search._obstacles_hp):
            hp[i, :len(h)] = h
            n_hp[i] = len(h)
        failed = 0
        try:
            cost, path, traj = search.run(debug=True)
        except Exception as ex:
            assert "No solution" in str(ex)
            failed, cost, path, traj = 1, np.nan, [], np.zeros((0, 3))
        seq = np.full(MAX_PATH, -1, np.int32)
        nodes = np.full((MAX_PATH, 3), np.nan)
        assert len(path) <= MAX_PATH
        for i, (a, b) in enumerate(zip(path[:-1], path[1:])):
            seq[i] = names.index(search._points_to_mp_names[a, b])
        for i, p in enumerate(path):
            nodes[i] = p
        tr = np.full(((MAX_PATH - 1) * 60, 3), np.nan)
        tr[:len(traj)] = traj
        ga = sc.goal_area
        rows["start"].append(sc.start); rows["goal"].append(sc.goal_point)
        rows["goal_box"].append([*ga.xy1, *ga.xy2]); rows["allowed"].append(sc.allowed_goal_theta_difference)
        rows["hp"].append(hp); rows["n_hp"].append(n_hp); rows["n_obs"].append(len(search._obstacles_hp))
        rows["cost"].append(cost); rows["n_path"].append(len(path)); rows["mp_seq"].append(seq); rows["nodes"].append(nodes)
        rows["n_expanded"].append(len(search.debug_data)); rows["traj"].append(tr); rows["n_traj"].append(len(traj))
        rows["start_pos"].append(s); rows["turn"].append(t); rows["failed"].append(failed)
        print(ci, s, t, "pert" if pert else "route", "cost", cost, "expanded", len(search.debug_data), "failed", failed,
              [names[k] for k in seq if k >= 0])
    out = {k: np.array(v) for k, v in rows.items()}
    out["cc_points"] = np.stack([search._mp_collision_points[n] for n in names])      # [9, 10, 3]
    out["radius"] = np.array(car.radius)
    out["circle_centers"] = np.array(car.circle_centers)
    path = os.path.join(REPO, "tests", "golden", "astar_intersection.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
