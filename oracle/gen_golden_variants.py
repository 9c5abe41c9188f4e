"""TEST INFRASTRUCTURE -- golden fixtures for the controller variants, recorded by running the
UNMODIFIED reference modules through oracle/refharness.py (build container only; needs /root/reference):

    python oracle/gen_golden_variants.py        # writes tests/golden/variant_*.npz

  tests/golden/variant_with_speed.npz
      main/lib/mpc_with_speed.py driven by the loop body of main/scenarios/mpc_intersection_new_ref.py:92-150
      (route (1,1), its two scripted cars, cutoff_idx passed to set_trajectory_fromarray instead of cutting the path),
      every MPC.step recorded, plus randomised single steps with random cutoff indices.
  tests/golden/variant_sensitivity.npz
      main/lib/mpc_sensitivity.py (re-reads ../config/mpc_config_sensitivity.json on every solve) for the 30
      (parameter, value) settings listed in main/scenarios/mpc_sensitivity_analysis_comulative.py:88-127, the first
      25 control steps of that script's loop (route (1,1), no obstacles) each.

As everywhere else the QP solve itself goes through oracle/cvxpy_shim (cvxpy -> ECOS is not installable here): parity
unpinned at that boundary, pinned everywhere else.
"""
import json
import math
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
sys.path.insert(0, HERE)
import refharness as rh  # noqa: E402
from gen_golden import DT, ROUTES, Recorder, record_step  # noqa: E402

SENSITIVITY_SWEEP = [  # main/scenarios/mpc_sensitivity_analysis_comulative.py:88-127
    ("w_perp", [0, 1, 10, 20, 50]),
    ("w_para", [0, 0.1, 1, 5, 10]),
    ("R", [[0, 0.01], [0.01, 0.01], [0.1, 0.01], [1, 0.01], [10, 0.01]]),
    ("R", [[0.1, 0.0], [0.1, 0.01], [0.1, 0.1], [0.1, 1], [0.1, 10]]),
    ("Rd", [[0, 1.0], [1, 1.0], [5, 1.0], [10, 1.0], [20, 1.0]]),
    ("Rd", [[10, 0], [10, 0.01], [10, 0.1], [10, 1], [10, 10]]),
]


def load_routes_raw():
    d = np.load(os.path.join(REPO, "mpc_for_av_at_intersection_b200", "data", "intersection_routes.npz"))
    return [d[f"route_{i}"].copy() for i in range(len(d["length"]))]


def with_speed(routes_raw, n_random=150, seed=77):
    rh.setup({})
    mod = rh.load_mpc("lib.mpc_with_speed")
    T = mod.T
    from lib.car_dimensions import BicycleModelDimensions
    from lib.collision_avoidance import check_collision_moving_cars, get_cutoff_curve_by_position_idx
    from lib.moving_obstacles import MovingObstacleTIntersection
    from lib.moving_obstacles_prediction import MovingObstaclesPrediction
    from lib.simulation import HistorySimulation, Simulation, State
    from lib.trajectories import calc_nearest_index_in_direction, resample_curve
    car = BicycleModelDimensions(skip_back_circle_collision_checking=False)
    rec = Recorder(T)
    MAX_SPEED = 30 / 3.6                                        # mpc_intersection_new_ref.py:32
    rid = ROUTES.index((1, 1))                                  # :37-38
    trajectory_full = routes_raw[rid].copy()
    dl = np.linalg.norm(trajectory_full[0, :2] - trajectory_full[1, :2])
    cv = np.full(trajectory_full[:, 1].shape, MAX_SPEED)        # :65
    mpc = mod.MPC(cx=trajectory_full[:, 0], cy=trajectory_full[:, 1], cv=cv, cyaw=trajectory_full[:, 2], dl=dl, dt=DT,
                  car_dimensions=car)
    state = State(x=trajectory_full[0, 0], y=trajectory_full[0, 1], yaw=trajectory_full[0, 2], v=0.0)
    simulation = HistorySimulation(car_dimensions=car, sample_time=DT, initial_state=state)
    obstacles = [MovingObstacleTIntersection(car, direction=1, offset=1., turning=False, speed=25 / 3.6, dt=DT),     # :42-44
                 MovingObstacleTIntersection(car, direction=-1, offset=4., turning=True, speed=25 / 3.6, dt=DT)]
    TIME_HORIZON, FRAME_WINDOW = 7., 20
    EXTRA = 4 * int(math.ceil(car.radius / dl))
    traj_agent_idx, tmp_trajectory = 0, None
    n_loop = 0
    for i in range(400):
        if mpc.is_goal(state):
            break
        if tmp_trajectory is None or np.any(tmp_trajectory[traj_agent_idx, :] != tmp_trajectory[-1, :]):
            traj_agent_idx = calc_nearest_index_in_direction(state, trajectory_full[:, 0], trajectory_full[:, 1],
                                                             start_index=traj_agent_idx, forward=True)
        trajectory_res = trajectory = trajectory_full[traj_agent_idx:]
        if state.v < Simulation.MAX_SPEED:
            resample_dl = np.zeros((trajectory_res.shape[0],)) + mod.MAX_ACCEL
            resample_dl = np.cumsum(resample_dl) + state.v
            resample_dl = DT * np.minimum(resample_dl, Simulation.MAX_SPEED)
            trajectory_res = resample_curve(trajectory_res, dl=resample_dl)
        else:
            trajectory_res = resample_curve(trajectory_res, dl=DT * Simulation.MAX_SPEED)
        trajs_mo = [np.vstack(MovingObstaclesPrediction(*o.get(), sample_time=DT, car_dimensions=car)
                              .state_prediction(TIME_HORIZON)).T for o in obstacles]
        collision_xy = check_collision_moving_cars(car, trajectory_res, trajectory, trajs_mo, frame_window=FRAME_WINDOW)
        cutoff_idx = 999                                        # mpc_intersection_new_ref.py:122-139
        if collision_xy is not None:
            cutoff_idx = get_cutoff_curve_by_position_idx(trajectory_full, collision_xy[0], collision_xy[1]) - EXTRA
            cutoff_idx = max(traj_agent_idx + 1, cutoff_idx)
        tmp_trajectory = trajectory_full
        mpc.set_trajectory_fromarray(tmp_trajectory, cutoff_idx=cutoff_idx)
        delta, acc, raised = record_step(rec, mpc, state, rid, 0, i, dict(cutoff_idx=cutoff_idx, kind=0))
        assert not raised
        for o in obstacles:
            o.step()
        state = simulation.step(a=acc, delta=delta, xref_deviation=mpc.get_current_xref_deviation())
        n_loop += 1
    h = simulation.history
    hist = np.array([h.x, h.y, h.yaw, h.v, h.delta, h.a, h.xref_deviation]).T
    print(f"with_speed closed loop: {n_loop} steps, cutoffs used: {sum(1 for c in rec.rows['cutoff_idx'] if c != 999)}")

    # randomised single steps: random route, state near the path, random warm start, random cutoff index
    rng = np.random.default_rng(seed)
    for k in range(n_random):
        rid = int(rng.integers(0, len(routes_raw)))
        full = routes_raw[rid].copy()
        dl = np.linalg.norm(full[0, :2] - full[1, :2])
        m = mod.MPC(cx=full[:, 0], cy=full[:, 1], cv=np.full(len(full), MAX_SPEED), cyaw=full[:, 2], dl=dl, dt=DT,
                    car_dimensions=car)
        j = int(rng.integers(0, len(full) - 60))
        cut = int(rng.choice([999, j + int(rng.integers(1, 120))]))
        m.set_trajectory_fromarray(full, cutoff_idx=cut)
        m.target_ind = max(j - 2, 0)
        lat = rng.normal(0, 0.3)
        yaw = full[j, 2]
        st = State(x=full[j, 0] - math.sin(yaw) * lat, y=full[j, 1] + math.cos(yaw) * lat,
                   yaw=yaw + rng.normal(0, 0.05), v=float(rng.uniform(0, 8.3)))
        if rng.random() < 0.7:
            m.oa = rng.uniform(-1.5, 1.5, T)
            m.odelta = np.clip(np.cumsum(rng.normal(0, 0.03, T)), -0.5, 0.5)
        record_step(rec, m, st, rid, 1, k, dict(cutoff_idx=cut, kind=1))
    rec.save(os.path.join(REPO, "tests", "golden", "variant_with_speed.npz"), history=hist, n_loop=np.array(n_loop),
             consts_json=np.array(json.dumps(dict(
                 T=int(mod.T), R=np.diag(mod.R).tolist(), Rd=np.diag(mod.Rd).tolist(), Q_v_yaw=np.diag(mod.Q_v_yaw).tolist(),
                 Qf_scaled=np.diag(mod.Qf).tolist(), GOAL_DIS=mod.GOAL_DIS, STOP_SPEED=mod.STOP_SPEED,
                 MAX_DSTEER=float(mod.MAX_DSTEER), MAX_ACCEL=mod.MAX_ACCEL, MAX_DECEL=mod.MAX_DECEL, MAX_SPEED=mod.MAX_SPEED))))


def sensitivity(routes_raw, steps=25):
    cfg0 = rh.setup({})
    mod = rh.load_mpc("lib.mpc_sensitivity")
    T = mod.T
    from lib.car_dimensions import BicycleModelDimensions
    from lib.simulation import HistorySimulation, State
    car = BicycleModelDimensions(skip_back_circle_collision_checking=False)
    rec = Recorder(T)
    rid = ROUTES.index((1, 1))                                  # mpc_sensitivity_analysis_comulative.py:134-135
    case = 0
    for name, values in SENSITIVITY_SWEEP:
        for val in values:
            config = dict(cfg0)
            config[name] = val
            with open("../config/mpc_config_sensitivity.json", "w") as f:       # :152-157
                json.dump(config, f, indent=4)
            trajectory_full = routes_raw[rid].copy()
            dl = np.linalg.norm(trajectory_full[0, :2] - trajectory_full[1, :2])
            mpc = mod.MPC(cx=trajectory_full[:, 0], cy=trajectory_full[:, 1], cyaw=trajectory_full[:, 2], dl=dl, dt=DT,
                          car_dimensions=car)
            state = State(x=trajectory_full[0, 0], y=trajectory_full[0, 1], yaw=trajectory_full[0, 2], v=0.0)
            simulation = HistorySimulation(car_dimensions=car, sample_time=DT, initial_state=state)
            w = [config["w_perp"], config["w_para"], *config["R"], *config["Rd"], *config["Q_v_yaw"], *config["Qf"]]
            for i in range(steps):
                if mpc.is_goal(state):
                    break
                mpc.set_trajectory_fromarray(trajectory_full)
                delta, acc, raised = record_step(rec, mpc, state, rid, case, i, dict(weights=np.array(w, float)))
                assert not raised
                state = simulation.step(a=acc, delta=delta, xref_deviation=mpc.get_current_xref_deviation())
            print(f"sensitivity case {case}: {name} = {val}")
            case += 1
    rec.save(os.path.join(REPO, "tests", "golden", "variant_sensitivity.npz"), n_cases=np.array(case))


if __name__ == "__main__":
    routes = load_routes_raw()
    with_speed(routes)
    sensitivity(routes)
