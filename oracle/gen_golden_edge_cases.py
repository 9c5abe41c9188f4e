"""TEST INFRASTRUCTURE -- tests/golden/edge_cases_T{10,13,20}.npz: MPC.step of the UNMODIFIED reference on a batch built to
exercise the exits of the step that ordinary problems never take, interleaved with ordinary problems so that neighbours in
the batch end differently (the two-problems-per-warp kernel runs neighbours in lockstep, mpcb_step_half.cuh):

  * the nearest-index Exception of main/lib/trajectories.py:120 ("something wrong"): about 1 in 12 000 random poses raises
    it (three nearest course points that are neither adjacent nor two apart) -- found here by screening 900 000 random poses
    with the C++ port (seeded), every one of them then confirmed by the reference itself (record_step asserts the message);
  * infeasible solves (v0 above MAX_SPEED / below MIN_SPEED: the constant rows x[2,0] <= MAX_SPEED of mpc.py:186-188)
    -> the reference keeps di and applies MAX_DECEL;
  * paths cut to one or two points, vehicles at the very end of the path, cold and warm starts;
  * an odd number of problems.

Same schema as tests/golden/random_T*.npz (oracle/gen_golden.py).  Run in the build container only:
    python oracle/gen_golden_edge_cases.py
"""
import math
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
sys.path.insert(0, HERE)
sys.path.insert(0, REPO)
import gen_golden as gg  # noqa: E402
import refharness as rh  # noqa: E402

DT = 0.2


def screen_raising_poses(routes_raw, n=900000, seed=7, keep=24):
    """(state xy/yaw/v, route, target_ind) triples on which the port reports the Exception class (status 3)."""
    import mpc_port
    from mpc_for_av_at_intersection_b200 import MpcConfig
    from mpc_for_av_at_intersection_b200.courses import CourseSet
    cs = CourseSet([r.copy() for r in routes_raw])
    cfgvec = mpc_port.pack_config(MpcConfig.from_json(T=10))
    rng = np.random.default_rng(seed)
    allxy = np.concatenate([r[:, :2] for r in routes_raw])
    lo, hi = allxy.min(0) - 5, allxy.max(0) + 5
    route = rng.integers(0, len(cs.length), n).astype(np.int32)
    nc = cs.length[route].astype(np.int32)
    tgt = (rng.uniform(0, 0.95, n) * nc).astype(np.int32)
    state = np.stack([rng.uniform(lo[0], hi[0], n), rng.uniform(lo[1], hi[1], n), rng.uniform(-np.pi, np.pi, n), rng.uniform(0, 6, n)], 1)
    out = mpc_port.step_batch(cfgvec, cs.packed, cs.dl, state, route, nc, np.full(n, 2.86), tgt.copy(), np.zeros(n, np.int32),
                              np.zeros((n, 10)), np.zeros((n, 10)), np.zeros((n, 2)))
    r = np.flatnonzero(out["status"] == 3)[:keep]
    print(f"screened {n} poses: {int((out['status'] == 3).sum())} raise, keeping {len(r)}")
    return state[r], route[r], tgt[r]


def edge_problems(T, routes_raw, raising, seed):
    cfg = rh.setup({"T": T})
    mpcmod = rh.load_mpc()
    from lib.car_dimensions import BicycleModelDimensions
    from lib.simulation import Simulation, State
    rng = np.random.default_rng(seed)
    rec = gg.Recorder(T)
    smoothed = []
    for r in routes_raw:
        r = r.copy()
        mpcmod.smooth_yaw(r[:, 2])
        smoothed.append(r)
    car = BicycleModelDimensions()
    rs, rr, rt = raising
    n_raise = 0
    N = 95                                              # odd on purpose
    for k in range(N):
        kind = k % 6                                    # neighbours differ: ordinary, raising, ordinary, infeasible, cut path, end of path
        rid = int(rng.integers(len(routes_raw)))
        if kind == 1 and n_raise < len(rs):
            rid = int(rr[n_raise])
        full = smoothed[rid]
        Nc = len(full)
        dl = np.linalg.norm(full[0, :2] - full[1, :2])
        i0 = int(rng.integers(0, Nc - 200))
        ncourse = Nc
        if kind == 5:
            i0 = int(rng.integers(Nc - 30, Nc))
        px, py, pyaw = full[i0]
        lat = rng.normal(0, 0.3)
        state = State(x=px - math.sin(pyaw) * lat, y=py + math.cos(pyaw) * lat, yaw=pyaw + rng.normal(0, 0.05), v=float(rng.uniform(0, 30 / 3.6)))
        target_in = max(0, i0 - int(rng.integers(0, 6)))
        if kind == 1 and n_raise < len(rs):
            state = State(x=float(rs[n_raise, 0]), y=float(rs[n_raise, 1]), yaw=float(rs[n_raise, 2]), v=float(rs[n_raise, 3]))
            target_in = int(rt[n_raise])
            n_raise += 1
        elif kind == 3:
            state.v = float(Simulation.MAX_SPEED + rng.uniform(0.5, 3.0)) if rng.random() < 0.7 else float(Simulation.MIN_SPEED - rng.uniform(0.5, 2.0))
        elif kind == 4:
            ncourse = min(Nc, max(1, target_in + int(rng.integers(-1, 3))))
        mpc = mpcmod.MPC(cx=full[:, 0], cy=full[:, 1], cyaw=full[:, 2].copy(), dl=dl, dt=DT, car_dimensions=car)
        mpc.set_trajectory_fromarray(full[:ncourse])
        mpc.target_ind = target_in
        if rng.random() < 0.6:
            mpc.oa = np.clip(rng.uniform(-3, 2) + np.cumsum(rng.normal(0, 0.3, T)), -10, 2)
            mpc.odelta = np.clip(np.cumsum(rng.normal(0, 0.03, T)), -0.7, 0.7)
            mpc.di, mpc.ai = float(mpc.odelta[0]), float(mpc.oa[0])
        gg.record_step(rec, mpc, state, rid, -1, k, dict(car=0, L=car.distance_back_to_front_wheel, dl=dl, kind=kind))
    rows = rec.rows
    raised, solved = np.array(rows["raised"]), np.array(rows["solved"])
    print(f"T={T}: {N} problems, {int(raised.sum())} raise, {int(((1 - raised) * (1 - solved)).sum())} fail, {int(solved.sum())} solve")
    assert raised.sum() >= 10 and ((1 - raised) * (1 - solved)).sum() >= 10
    rec.save(os.path.join(REPO, "tests", "golden", f"edge_cases_T{T}.npz"), config_json=np.array(str(cfg)))


def main():
    rh.setup()
    d = np.load(os.path.join(REPO, "mpc_for_av_at_intersection_b200", "data", "intersection_routes.npz"))
    routes_raw = [d[f"route_{i}"].copy() for i in range(len(d["length"]))]
    raising = screen_raising_poses(routes_raw)
    for T in (10, 13, 20):
        edge_problems(T, routes_raw, raising, seed=4000 + T)


if __name__ == "__main__":
    main()
