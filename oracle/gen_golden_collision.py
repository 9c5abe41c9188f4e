"""TEST INFRASTRUCTURE -- record golden vectors for the coupling step (row A9) by calling the UNMODIFIED
reference functions (build container only; needs /root/reference):

    python oracle/gen_golden_collision.py        # writes tests/golden/collision_random.npz

Per sample: a route, the ego's index/speed on it, 1..7 obstacles in MovingObstacle*.get() layout
(x, y, v, yaw, a, steer); outputs of resample_curve (length), check_collision_moving_cars (x, y, idx or none) and
the cutoff index exactly as main/scenarios/mpc_intersection.py:108-136 composes them."""
import math
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
sys.path.insert(0, HERE)
import refharness as rh  # noqa: E402

DT, TIME_HORIZON, FRAME_WINDOW, MAX_OBS = 0.2, 7.0, 20, 7


def main(n=360, seed=4242):
    rh.setup()
    mpcmod = rh.load_mpc()
    from lib.car_dimensions import BicycleModelDimensions, PriusDimensions
    from lib.collision_avoidance import check_collision_moving_cars, get_cutoff_curve_by_position_idx
    from lib.moving_obstacles_prediction import MovingObstaclesPrediction
    from lib.simulation import Simulation
    from lib.trajectories import resample_curve
    d = np.load(os.path.join(REPO, "mpc_for_av_at_intersection_b200", "data", "intersection_routes.npz"))
    routes = [d[f"route_{i}"].copy() for i in range(len(d["length"]))]
    for r in routes:
        mpcmod.smooth_yaw(r[:, 2])
    cars = [BicycleModelDimensions(), PriusDimensions()]
    rng = np.random.default_rng(seed)
    rec = {k: [] for k in ("route", "car", "agent_idx", "v", "n_obs", "obs", "n_res", "collided", "coll_x", "coll_y",
                           "coll_idx", "cutoff_idx")}
    for k in range(n):
        rid = int(rng.integers(len(routes)))
        full = routes[rid]
        Nc = len(full)
        car_id = int(rng.random() < 0.2)
        car = cars[car_id]
        dl = np.linalg.norm(full[0, :2] - full[1, :2])
        EXTRA = 4 * int(math.ceil(car.radius / dl))
        u = rng.random()
        agent_idx = int(rng.integers(0, Nc)) if u < 0.85 else int(rng.integers(Nc - 3, Nc))
        v = float(rng.uniform(0, 30 / 3.6))
        if rng.random() < 0.1:
            v = 30 / 3.6 if rng.random() < 0.5 else 0.0
        n_obs = int(rng.integers(1, MAX_OBS + 1))
        obs = np.full((MAX_OBS, 6), np.nan)
        for o in range(n_obs):
            if rng.random() < 0.35:     # near the ego's future path
                j = min(Nc - 1, agent_idx + int(rng.integers(0, 420)))
                px, py = full[j, 0] + rng.normal(0, 3.0), full[j, 1] + rng.normal(0, 3.0)
            else:
                px, py = rng.uniform(-35, 35), rng.uniform(-35, 35)
            yaw = rng.uniform(-math.pi, 2 * math.pi)
            vo = float(rng.uniform(0, 9)) if rng.random() < 0.8 else 0.0
            a = 0.0 if rng.random() < 0.6 else float(rng.uniform(-1.5, 1.5))
            st = float(rng.choice([0.0, 0.0, 0.19, -0.38, rng.uniform(-0.5, 0.5)]))
            obs[o] = (px, py, vo, yaw, a, st)
        # ---- the reference's loop body, main/scenarios/mpc_intersection.py:108-136 --------------------------
        trajectory_res = trajectory = full[agent_idx:]
        if v < Simulation.MAX_SPEED:
            resample_dl = np.zeros((trajectory_res.shape[0],)) + mpcmod.MAX_ACCEL
            resample_dl = np.cumsum(resample_dl) + v
            resample_dl = DT * np.minimum(resample_dl, Simulation.MAX_SPEED)
            trajectory_res = resample_curve(trajectory_res, dl=resample_dl)
        else:
            trajectory_res = resample_curve(trajectory_res, dl=DT * Simulation.MAX_SPEED)
        trajs = [np.vstack(MovingObstaclesPrediction(*obs[o], sample_time=DT, car_dimensions=car)
                           .state_prediction(TIME_HORIZON)).T for o in range(n_obs)]
        hit = check_collision_moving_cars(car, trajectory_res, trajectory, trajs, frame_window=FRAME_WINDOW)
        if hit is not None:
            cutoff = get_cutoff_curve_by_position_idx(full, hit[0], hit[1]) - EXTRA
            cutoff = max(agent_idx + 1, cutoff)
        else:
            cutoff = Nc
        for key, val in (("route", rid), ("car", car_id), ("agent_idx", agent_idx), ("v", v), ("n_obs", n_obs),
                         ("obs", obs), ("n_res", len(trajectory_res)), ("collided", int(hit is not None)),
                         ("coll_x", hit[0] if hit else np.nan), ("coll_y", hit[1] if hit else np.nan),
                         ("coll_idx", hit[2] if hit else -1), ("cutoff_idx", cutoff)):
            rec[key].append(val)
    out = {k: np.array(v) for k, v in rec.items()}
    path = os.path.join(REPO, "tests", "golden", "collision_random.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, "collided fraction", out["collided"].mean(), "n_res range", out["n_res"].min(), out["n_res"].max())


if __name__ == "__main__":
    main()
