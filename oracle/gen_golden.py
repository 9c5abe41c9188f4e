"""TEST INFRASTRUCTURE -- generate the committed golden fixtures by running the UNMODIFIED reference.

Run in the build container only (needs /root/reference):

    python oracle/gen_golden.py            # writes tests/golden/*.npz and the route library

What is recorded (all float64 / int64, reference functions called as-is through
oracle/refharness.py; the only non-reference arithmetic is the QP solve, which the
reference delegates to cvxpy -> ECOS and which oracle/qp_ipm.py replaces -- parity
unpinned at that boundary, see oracle/README.md):

  mpc_for_av_at_intersection_b200/data/intersection_routes.npz
      the 12 (start_pos, turn) A* routes of main/envs/intersection.py planned by
      main/lib/motion_primitive_search_modified.py (raw, before smooth_yaw)
  tests/golden/closed_loop_T{13,20}.npz
      per-step inputs/outputs of MPC.step for the 12 routes without obstacles and for
      route (4,1) with the two scripted obstacles of main/scenarios/mpc_intersection.py:31-45
      (loop body restated from :95-159, plotting dropped)
  tests/golden/random_T{13,20}.npz
      randomised single MPC.step problems (perturbed states, warm starts, cut paths)
"""
import math
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
sys.path.insert(0, HERE)
import refharness as rh  # noqa: E402

ROUTES = [(s, t) for s in (1, 2, 3, 4) for t in (1, 2, 3)]
DT = 0.2


class Recorder:
    def __init__(self, T):
        self.T = T
        self.rows = {}

    def add(self, **kw):
        for k, v in kw.items():
            self.rows.setdefault(k, []).append(np.asarray(v))

    def save(self, path, **extra):
        out = {k: np.stack(v) for k, v in self.rows.items()}
        out.update(extra)
        np.savez_compressed(path, **out)
        print("wrote", path, {k: v.shape for k, v in out.items() if hasattr(v, "shape")})


def record_step(rec, mpc, state, route_id, loop_id, step, extra=None):
    """Call the reference MPC.step and record inputs + outputs."""
    T = rec.T
    nan = np.full(T, np.nan)
    nan1 = np.full(T + 1, np.nan)
    inp = dict(
        state=[state.x, state.y, state.yaw, state.v], route=route_id, loop=loop_id, step=step,
        ncourse=len(mpc.cx), target_in=mpc.target_ind,
        has_warm=int(mpc.oa is not None and mpc.odelta is not None),
        oa_in=nan if mpc.oa is None else np.asarray(mpc.oa, float),
        od_in=nan if mpc.odelta is None else np.asarray(mpc.odelta, float),
        di_in=mpc.di, ai_in=mpc.ai)
    raised = 0
    try:
        di, ai = mpc.step(state)
    except Exception as ex:  # trajectories.py:120
        assert "something wrong" in str(ex), ex
        raised = 1
        di, ai = mpc.di, mpc.ai
    ok = (not raised) and mpc.odelta is not None
    xref = mpc.xref if (not raised) else np.full((4, T + 1), np.nan)
    out = dict(
        raised=raised, solved=int(ok), di=di, ai=ai, target_out=mpc.target_ind,
        oa=np.asarray(mpc.oa, float) if ok else nan, od=np.asarray(mpc.odelta, float) if ok else nan,
        ox=mpc.ox if ok else nan1, oy=mpc.oy if ok else nan1, ov=mpc.ov if ok else nan1,
        oyaw=mpc.oyaw if ok else nan1, xref=xref,
        xref_dev=mpc.get_current_xref_deviation() if (ok and mpc.target_ind < len(mpc.cx)) else np.nan)  # IndexError in the reference otherwise
    rec.add(**inp, **out, **(extra or {}))
    return di, ai, raised


def closed_loops(T, routes_raw, cases=None, max_steps=400):
    cfg = rh.setup({"T": T})
    mpcmod = rh.load_mpc()
    from lib.car_dimensions import BicycleModelDimensions
    from lib.collision_avoidance import check_collision_moving_cars, get_cutoff_curve_by_position_idx
    from lib.moving_obstacles import MovingObstacleTIntersection
    from lib.moving_obstacles_prediction import MovingObstaclesPrediction
    from lib.simulation import HistorySimulation, Simulation, State
    from lib.trajectories import calc_nearest_index_in_direction, resample_curve

    car = BicycleModelDimensions(skip_back_circle_collision_checking=False)
    rec = Recorder(T)
    if cases is None:
        cases = [(rid, False) for rid in range(len(ROUTES))] + [(ROUTES.index((4, 1)), True)]
    loop_len = []
    for loop_id, (rid, with_obs) in enumerate(cases):
        trajectory_full = routes_raw[rid].copy()
        dl = np.linalg.norm(trajectory_full[0, :2] - trajectory_full[1, :2])
        mpc = mpcmod.MPC(cx=trajectory_full[:, 0], cy=trajectory_full[:, 1], cyaw=trajectory_full[:, 2],
                         dl=dl, dt=DT, car_dimensions=car)
        state = State(x=trajectory_full[0, 0], y=trajectory_full[0, 1], yaw=trajectory_full[0, 2], v=0.0)
        simulation = HistorySimulation(car_dimensions=car, sample_time=DT, initial_state=state)
        obstacles = []
        if with_obs:  # mpc_intersection.py:41-44
            obstacles = [
                MovingObstacleTIntersection(car, direction=1, offset=2., turning=False, speed=25 / 3.6, dt=DT),
                MovingObstacleTIntersection(car, direction=-1, offset=4., turning=True, speed=25 / 3.6, dt=DT)]
        TIME_HORIZON, FRAME_WINDOW = 7., 20
        EXTRA = 4 * int(math.ceil(car.radius / dl))
        traj_agent_idx = 0
        tmp_trajectory = None
        for i in range(max_steps):
            if mpc.is_goal(state):
                break
            # mpc_intersection.py:103-139
            if tmp_trajectory is None or np.any(tmp_trajectory[traj_agent_idx, :] != tmp_trajectory[-1, :]):
                traj_agent_idx = calc_nearest_index_in_direction(
                    state, trajectory_full[:, 0], trajectory_full[:, 1], start_index=traj_agent_idx, forward=True)
            trajectory_res = trajectory = trajectory_full[traj_agent_idx:]
            if state.v < Simulation.MAX_SPEED:
                resample_dl = np.zeros((trajectory_res.shape[0],)) + mpcmod.MAX_ACCEL
                resample_dl = np.cumsum(resample_dl) + state.v
                resample_dl = DT * np.minimum(resample_dl, Simulation.MAX_SPEED)
                trajectory_res = resample_curve(trajectory_res, dl=resample_dl)
            else:
                trajectory_res = resample_curve(trajectory_res, dl=DT * Simulation.MAX_SPEED)
            obs_states = np.array([o.get() for o in obstacles], float).reshape(-1, 6)
            trajs_mo = [np.vstack(MovingObstaclesPrediction(*o.get(), sample_time=DT, car_dimensions=car)
                                  .state_prediction(TIME_HORIZON)).T for o in obstacles]
            collision_xy = check_collision_moving_cars(car, trajectory_res, trajectory, trajs_mo,
                                                       frame_window=FRAME_WINDOW)
            if collision_xy is not None:
                cutoff_idx = get_cutoff_curve_by_position_idx(trajectory_full, collision_xy[0],
                                                              collision_xy[1]) - EXTRA
                cutoff_idx = max(traj_agent_idx + 1, cutoff_idx)
                tmp_trajectory = trajectory_full[:cutoff_idx]
            else:
                cutoff_idx = len(trajectory_full)
                tmp_trajectory = trajectory_full
            mpc.set_trajectory_fromarray(tmp_trajectory)
            extra = dict(
                agent_idx=traj_agent_idx, n_res=len(trajectory_res), collided=int(collision_xy is not None),
                coll_x=collision_xy[0] if collision_xy is not None else np.nan,
                coll_y=collision_xy[1] if collision_xy is not None else np.nan,
                coll_idx=collision_xy[2] if collision_xy is not None else -1, cutoff_idx=cutoff_idx,
                obs=np.vstack([obs_states, np.full((2 - len(obs_states), 6), np.nan)]))
            delta, acc, raised = record_step(rec, mpc, state, rid, loop_id, i, extra)
            assert not raised
            for o in obstacles:
                o.step()
            state = simulation.step(a=acc, delta=delta, xref_deviation=mpc.get_current_xref_deviation())
        loop_len.append(i)
        h = simulation.history
        rec.rows.setdefault("_hist", []).append(np.array([h.x, h.y, h.yaw, h.v, h.delta, h.a, h.xref_deviation]).T)
        print(f"T={T} loop {loop_id} route {ROUTES[rid]} obstacles={with_obs}: {i} steps")
    hist = rec.rows.pop("_hist")
    hist_flat = np.concatenate(hist)
    hist_off = np.cumsum([0] + [len(h) for h in hist])
    rec.save(os.path.join(REPO, "tests", "golden", f"closed_loop_T{T}.npz"),
             loop_len=np.array(loop_len), loop_route=np.array([c[0] for c in cases]),
             loop_obstacles=np.array([int(c[1]) for c in cases]), history=hist_flat, history_off=hist_off,
             config_json=np.array(str(cfg)))


def random_problems(T, routes_raw, n, seed):
    cfg = rh.setup({"T": T})
    mpcmod = rh.load_mpc()
    from lib.car_dimensions import BicycleModelDimensions, PriusDimensions
    from lib.simulation import State
    rng = np.random.default_rng(seed)
    rec = Recorder(T)
    smoothed = []
    for r in routes_raw:
        r = r.copy()
        mpcmod.smooth_yaw(r[:, 2])
        smoothed.append(r)
    cars = [BicycleModelDimensions(), PriusDimensions()]
    for k in range(n):
        rid = int(rng.integers(len(routes_raw)))
        full = smoothed[rid]
        Nc = len(full)
        car_id = int(rng.random() < 0.25)
        car = cars[car_id]
        dl = np.linalg.norm(full[0, :2] - full[1, :2])
        kind = rng.random()
        i0 = int(rng.integers(0, Nc - 200)) if kind < 0.75 else int(rng.integers(Nc - 200, Nc))
        px, py, pyaw = full[i0]
        lat = rng.normal(0, 0.3)
        state = State(x=px - math.sin(pyaw) * lat + rng.normal(0, 0.02),
                      y=py + math.cos(pyaw) * lat + rng.normal(0, 0.02),
                      yaw=pyaw + rng.normal(0, 0.05), v=float(rng.uniform(0, 30 / 3.6)))
        if rng.random() < 0.1:
            state.v = 30 / 3.6 if rng.random() < 0.5 else 0.0
        target_in = max(0, i0 - int(rng.integers(0, 6)))
        c = rng.random()
        if c < 0.7:
            ncourse = Nc
        elif c < 0.9:
            ncourse = int(rng.integers(min(target_in + 3, Nc), Nc + 1))
        else:
            ncourse = min(Nc, max(1, target_in + int(rng.integers(-1, 3))))
        mpc = mpcmod.MPC(cx=full[:, 0], cy=full[:, 1], cyaw=full[:, 2].copy(), dl=dl, dt=DT, car_dimensions=car)
        mpc.set_trajectory_fromarray(full[:ncourse])
        mpc.target_ind = target_in
        if rng.random() < 0.6:
            base_a = rng.uniform(-3, 2)
            mpc.oa = np.clip(base_a + np.cumsum(rng.normal(0, 0.3, T)), -10, 2)
            mpc.odelta = np.clip(np.cumsum(rng.normal(0, 0.03, T)), -0.7, 0.7)
            mpc.di, mpc.ai = float(mpc.odelta[0]), float(mpc.oa[0])
        record_step(rec, mpc, state, rid, -1, k, dict(car=car_id, L=car.distance_back_to_front_wheel, dl=dl))
    rec.save(os.path.join(REPO, "tests", "golden", f"random_T{T}.npz"), config_json=np.array(str(cfg)))


def main():
    os.makedirs(os.path.join(REPO, "tests", "golden"), exist_ok=True)
    data_dir = os.path.join(REPO, "mpc_for_av_at_intersection_b200", "data")
    os.makedirs(data_dir, exist_ok=True)
    rh.setup()
    routes_raw, costs = [], []
    for (s, t) in ROUTES:
        cost, path, traj = rh.plan_route(s, t)
        routes_raw.append(np.asarray(traj, float))
        costs.append(cost)
        print("route", (s, t), "cost", cost, "points", traj.shape)
    np.savez_compressed(
        os.path.join(data_dir, "intersection_routes.npz"),
        start_turn=np.array(ROUTES), cost=np.array(costs), length=np.array([len(r) for r in routes_raw]),
        **{f"route_{i}": r for i, r in enumerate(routes_raw)})
    for T in (13, 20):
        closed_loops(T, routes_raw)
        random_problems(T, routes_raw, n=400, seed=1000 + T)


def main_long():
    """Long horizons of BASELINE.json configs[4] (T = 40, 80): fewer cases, the oracle QP has 6T unknowns.
    python oracle/gen_golden.py --long"""
    rh.setup()
    z = np.load(os.path.join(REPO, "mpc_for_av_at_intersection_b200", "data", "intersection_routes.npz"))
    routes_raw = [z[f"route_{i}"] for i in range(len(ROUTES))]
    random_problems(40, routes_raw, n=96, seed=1040)
    random_problems(80, routes_raw, n=48, seed=1080)
    # one straight route and the scripted-obstacle left turn, truncated (every step is an independent check)
    closed_loops(40, routes_raw, cases=[(ROUTES.index((2, 2)), False), (ROUTES.index((4, 1)), True)], max_steps=60)
    closed_loops(80, routes_raw, cases=[(ROUTES.index((4, 1)), True)], max_steps=30)


if __name__ == "__main__":
    if "--long" in sys.argv:
        main_long()
    else:
        main()
