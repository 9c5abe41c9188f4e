"""MEASUREMENT INFRASTRUCTURE (build container only; needs /root/reference) -- wall-clock of the UNMODIFIED reference loop
body of main/scenarios/mpc_intersection.py:95-159 (route (4, 1), its two scripted cars, :41-44), the two figures the script
itself prints and never records (:96,145-147,161-167): `mpc.step` alone and the whole loop body, one process = one core.
The QP solve goes through oracle/cvxpy_shim (dense numpy interior point + polish), NOT cvxpy -> ECOS, which cannot be
installed here: these numbers are context for BASELINE.md, not a reference-published baseline.

usage: python oracle/time_reference_loop.py [T ...]"""
import math
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import refharness as rh  # noqa: E402
from gen_golden import DT, ROUTES  # noqa: E402

REPO = os.path.dirname(HERE)


def run(T):
    rh.setup({"T": T})
    mod = rh.load_mpc()
    from lib.car_dimensions import BicycleModelDimensions
    from lib.collision_avoidance import check_collision_moving_cars, get_cutoff_curve_by_position_idx
    from lib.moving_obstacles import MovingObstacleTIntersection
    from lib.moving_obstacles_prediction import MovingObstaclesPrediction
    from lib.simulation import HistorySimulation, Simulation, State
    from lib.trajectories import calc_nearest_index_in_direction, resample_curve
    car = BicycleModelDimensions(skip_back_circle_collision_checking=False)
    d = np.load(os.path.join(REPO, "mpc_for_av_at_intersection_b200", "data", "intersection_routes.npz"))
    full = d[f"route_{ROUTES.index((4, 1))}"].copy()
    dl = np.linalg.norm(full[0, :2] - full[1, :2])
    mpc = mod.MPC(cx=full[:, 0], cy=full[:, 1], cyaw=full[:, 2], dl=dl, dt=DT, car_dimensions=car)
    state = State(x=full[0, 0], y=full[0, 1], yaw=full[0, 2], v=0.0)
    sim = HistorySimulation(car_dimensions=car, sample_time=DT, initial_state=state)
    obstacles = [MovingObstacleTIntersection(car, direction=1, offset=2., turning=False, speed=25 / 3.6, dt=DT),
                 MovingObstacleTIntersection(car, direction=-1, offset=4., turning=True, speed=25 / 3.6, dt=DT)]
    EXTRA = 4 * int(math.ceil(car.radius / dl))
    idx, tmp = 0, None
    t_step, t_body = [], []
    for i in range(400):
        t0 = time.perf_counter()
        if mpc.is_goal(state):
            break
        if tmp is None or np.any(tmp[idx, :] != tmp[-1, :]):
            idx = calc_nearest_index_in_direction(state, full[:, 0], full[:, 1], start_index=idx, forward=True)
        res = traj = full[idx:]
        if state.v < Simulation.MAX_SPEED:
            rdl = DT * np.minimum(np.cumsum(np.zeros((res.shape[0],)) + mod.MAX_ACCEL) + state.v, Simulation.MAX_SPEED)
            res = resample_curve(res, dl=rdl)
        else:
            res = resample_curve(res, dl=DT * Simulation.MAX_SPEED)
        tm = [np.vstack(MovingObstaclesPrediction(*o.get(), sample_time=DT, car_dimensions=car).state_prediction(7.)).T for o in obstacles]
        hit = check_collision_moving_cars(car, res, traj, tm, frame_window=20)
        if hit is not None:
            tmp = full[:max(idx + 1, get_cutoff_curve_by_position_idx(full, hit[0], hit[1]) - EXTRA)]
        else:
            tmp = full
        mpc.set_trajectory_fromarray(tmp)
        t1 = time.perf_counter()
        delta, acc = mpc.step(state)
        t2 = time.perf_counter()
        for o in obstacles:
            o.step()
        state = sim.step(a=acc, delta=delta, xref_deviation=mpc.get_current_xref_deviation())
        t_step.append(t2 - t1)
        t_body.append(time.perf_counter() - t0)
    s, b = np.array(t_step) * 1e3, np.array(t_body) * 1e3
    print(f"T={T}: {len(s)} control steps; mpc.step mean {s.mean():.1f} ms p50 {np.percentile(s, 50):.1f} ms -> {1e3 / s.mean():.0f} solves/s on one core; "
          f"whole loop body mean {b.mean():.1f} ms p50 {np.percentile(b, 50):.1f} ms (non-MPC part {b.mean() - s.mean():.1f} ms)")


if __name__ == "__main__":
    for T in [int(a) for a in sys.argv[1:]] or [13, 20]:
        run(T)
