"""TEST INFRASTRUCTURE -- golden fixtures for the jerk controller variant, recorded by running the
UNMODIFIED main/lib/mpc_jerk.py through oracle/refharness.py (build container only; needs /root/reference):

    python oracle/gen_golden_jerk.py        # writes tests/golden/variant_jerk.npz

main/lib/mpc_jerk.py: NX = 5 (fifth state = the integral of the acceleration input with a FREE initial value, :16,67-78),
hard-coded weights (Rd = (0.3, 1), w_perp = 10, w_para = 1, :21-27,167-171), MAX_DECEL = -5 (:39), a jerk penalty on the
fifth state's increments (:190).  No script of the reference imports it, so the loop here is the body of
main/scenarios/mpc_intersection.py:95-159 without obstacles on two routes, plus randomised single steps.
The QP solve goes through oracle/cvxpy_shim (parity unpinned at the cvxpy -> ECOS boundary, pinned everywhere else)."""
import json
import math
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
sys.path.insert(0, HERE)
import refharness as rh  # noqa: E402
from gen_golden import DT, ROUTES, Recorder  # noqa: E402


def record_step(rec, mpc, state, route_id, loop_id, step, extra=None):
    T = rec.T
    nan, nan1 = np.full(T, np.nan), np.full(T + 1, np.nan)
    inp = dict(state=[state.x, state.y, state.yaw, state.v], route=route_id, loop=loop_id, step=step, ncourse=len(mpc.cx),
               target_in=mpc.target_ind, has_warm=int(mpc.oa is not None and mpc.odelta is not None),
               oa_in=nan if mpc.oa is None else np.asarray(mpc.oa, float),
               od_in=nan if mpc.odelta is None else np.asarray(mpc.odelta, float), di_in=mpc.di, ai_in=mpc.ai)
    raised = 0
    try:
        di, ai = mpc.step(state)
    except Exception as ex:
        assert "something wrong" in str(ex), ex
        raised = 1
        di, ai = mpc.di, mpc.ai
    ok = (not raised) and mpc.odelta is not None
    out = dict(raised=raised, solved=int(ok), di=di, ai=ai, target_out=mpc.target_ind,
               oa=np.asarray(mpc.oa, float) if ok else nan, od=np.asarray(mpc.odelta, float) if ok else nan,
               ox=mpc.ox if ok else nan1, oy=mpc.oy if ok else nan1, ov=mpc.ov if ok else nan1, oyaw=mpc.oyaw if ok else nan1,
               xref=np.asarray(mpc.xref)[:4] if not raised else np.full((4, T + 1), np.nan))
    rec.add(**inp, **out, **(extra or {}))
    return di, ai, raised


def main(n_random=160, seed=91):
    rh.setup({})
    mod = rh.load_mpc("lib.mpc_jerk")
    T = mod.T
    from lib.car_dimensions import BicycleModelDimensions
    from lib.simulation import HistorySimulation, State
    car = BicycleModelDimensions(skip_back_circle_collision_checking=False)
    d = np.load(os.path.join(REPO, "mpc_for_av_at_intersection_b200", "data", "intersection_routes.npz"))
    routes_raw = [d[f"route_{i}"].copy() for i in range(len(d["length"]))]
    rec = Recorder(T)
    hists, loop_len = [], []
    for loop_id, key in enumerate([(4, 1), (1, 2)]):
        rid = ROUTES.index(key)
        full = routes_raw[rid].copy()
        dl = np.linalg.norm(full[0, :2] - full[1, :2])
        mpc = mod.MPC(cx=full[:, 0], cy=full[:, 1], cyaw=full[:, 2], dl=dl, dt=DT, car_dimensions=car)
        state = State(x=full[0, 0], y=full[0, 1], yaw=full[0, 2], v=0.0)
        sim = HistorySimulation(car_dimensions=car, sample_time=DT, initial_state=state)
        n = 0
        for i in range(400):
            if mpc.is_goal(state):
                break
            delta, acc, raised = record_step(rec, mpc, state, rid, loop_id, i, dict(kind=0))
            assert not raised
            state = sim.step(a=acc, delta=delta, xref_deviation=mpc.get_current_xref_deviation())
            n += 1
        h = sim.history
        hists.append(np.array([h.x, h.y, h.yaw, h.v, h.delta, h.a, h.xref_deviation]).T)
        loop_len.append(n)
        print(f"jerk closed loop route {key}: {n} steps")
    rng = np.random.default_rng(seed)
    for k in range(n_random):
        rid = int(rng.integers(0, len(routes_raw)))
        full = routes_raw[rid].copy()
        dl = np.linalg.norm(full[0, :2] - full[1, :2])
        m = mod.MPC(cx=full[:, 0], cy=full[:, 1], cyaw=full[:, 2], dl=dl, dt=DT, car_dimensions=car)
        j = int(rng.integers(0, len(full) - 40))
        if rng.random() < 0.25:
            m.set_trajectory_fromarray(full[: j + int(rng.integers(1, 150))])
        m.target_ind = max(j - 2, 0)
        lat, yaw = rng.normal(0, 0.3), full[j, 2]
        st = State(x=full[j, 0] - math.sin(yaw) * lat, y=full[j, 1] + math.cos(yaw) * lat, yaw=yaw + rng.normal(0, 0.05),
                   v=float(rng.uniform(0, 8.3)))
        if rng.random() < 0.7:
            m.oa = rng.uniform(-1.5, 1.5, T)
            m.odelta = np.clip(np.cumsum(rng.normal(0, 0.03, T)), -0.5, 0.5)
        record_step(rec, m, st, rid, 9, k, dict(kind=1))
    rec.save(os.path.join(REPO, "tests", "golden", "variant_jerk.npz"), history=np.concatenate(hists), loop_len=np.array(loop_len),
             consts_json=np.array(json.dumps(dict(
                 T=int(mod.T), R=np.diag(mod.R).tolist(), Rd=np.diag(mod.Rd).tolist(), Q_v_yaw=np.diag(mod.Q_v_yaw).tolist(),
                 Qf_scaled=np.diag(mod.Qf).tolist(), GOAL_DIS=mod.GOAL_DIS, STOP_SPEED=mod.STOP_SPEED, MAX_DSTEER=float(mod.MAX_DSTEER),
                 MAX_ACCEL=mod.MAX_ACCEL, MAX_DECEL=mod.MAX_DECEL, jerk_penalty_weight=mod.jerk_penalty_weight))))


if __name__ == "__main__":
    main()
