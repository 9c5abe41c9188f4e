"""SURVEY 8f N2 measurement: B randomised planning problems (all 12 (start lane, turn) pairs of the junction, start pose
perturbed) in ONE mpcb_plan_routes_host call, next to the CPU oracle (the reference planner's algorithm in Python, like
the reference) on a bounded sample.  usage: python scripts/bench_planner.py [B]  ->  one JSON line"""
import dataclasses, json, os, sys, time
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO); sys.path.insert(0, os.path.join(REPO, "oracle"))
from mpc_for_av_at_intersection_b200 import planner
from mpc_for_av_at_intersection_b200.compat.car_dimensions import BicycleModelDimensions

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
car = BicycleModelDimensions(skip_back_circle_collision_checking=False)
lib = planner.PrimitiveLibrary(planner.load_motion_primitives("bicycle_model"), car)
rng = np.random.default_rng(4242)
base = {(s, t): planner.intersection(turn_indicator=t, start_pos=s) for s in (1, 2, 3, 4) for t in (1, 2, 3)}
scs = []
for _ in range(B):
    sc = base[(int(rng.integers(1, 5)), int(rng.integers(1, 4)))]
    x, y, th = sc.start
    lon, lat, dth = rng.uniform(-1.0, 3.0), rng.normal(0, 0.15), rng.normal(0, 0.015)
    scs.append(dataclasses.replace(sc, start=(x + np.cos(th) * lon - np.sin(th) * lat, y + np.sin(th) * lon + np.cos(th) * lat, th + dth)))
planner.plan_batch(scs[:64], lib, margin=car.radius, max_open=32768, max_closed=16384)        # warm-up: context, module load
passes = []
t0 = time.perf_counter()
r = planner.plan_batch(scs, lib, margin=car.radius, max_open=32768, max_closed=16384, grow=0)
passes.append(dict(searches=B, max_open=32768, max_closed=16384, wall_s=time.perf_counter() - t0, overflow=int((r.status == planner.PLAN_OVERFLOW).sum())))
# the overflowed searches again, alone, at 4x the capacities (what plan_batch(grow=...) does), timed per pass
mo, mc = 32768, 16384
for _ in range(int(os.environ.get("PLANNER_GROW", "2"))):
    ov = np.flatnonzero(r.status == planner.PLAN_OVERFLOW)
    if len(ov) == 0:
        break
    mo, mc = 4 * mo, 4 * mc
    t1 = time.perf_counter()
    sub = planner.plan_batch([scs[b] for b in ov], lib, margin=car.radius, max_open=mo, max_closed=mc, grow=0)
    for name in ("status", "cost", "n_path", "mp_seq", "nodes", "n_expanded", "n_pushed", "traj", "n_traj"):
        getattr(r, name)[ov] = getattr(sub, name)
    passes.append(dict(searches=len(ov), max_open=mo, max_closed=mc, wall_s=time.perf_counter() - t1, overflow=int((sub.status == planner.PLAN_OVERFLOW).sum())))
t_gpu = time.perf_counter() - t0
ok = r.status == planner.PLAN_OK
# CPU: the oracle on a bounded sample (searches the GPU finished within 3000 expansions, so the sample stays short)
import astar_oracle as ao
sample = [k for k in range(B) if ok[k] and r.n_expanded[k] <= 3000][:24]
t0 = time.perf_counter()
exp_cpu = 0
for k in sample:
    sc = scs[k]
    o = ao.AStarOracle(sc.start, sc.goal_point, [*sc.goal_area.xy1, *sc.goal_area.xy2], sc.allowed_goal_theta_difference,
                       [ob.to_convex(margin=car.radius) for ob in sc.obstacles], lib.points, lib.n_points, lib.length, lib.cc)
    res = o.run()
    assert res is not None and res[2] == list(r.mp_seq[k, :r.n_path[k] - 1]) and o.n_expanded == r.n_expanded[k]
    exp_cpu += o.n_expanded
t_cpu = time.perf_counter() - t0
print(json.dumps({"metric": "planner_searches_per_sec", "workload": f"{B} A* searches on the reference junction, 9 primitives, 28 obstacles, randomised start poses; one synchronous call incl. workspace allocation and all copies",
                  "value": B / t_gpu, "wall_s": t_gpu, "expanded_nodes_total": int(r.n_expanded.sum()), "expanded_nodes_per_sec": float(r.n_expanded.sum() / t_gpu),
                  "expanded_p50": int(np.median(r.n_expanded)), "expanded_p99": int(np.quantile(r.n_expanded, 0.99)), "expanded_max": int(r.n_expanded.max()),
                  "passes": passes, "status_counts": {"ok": int(ok.sum()), "no_solution": int((r.status == 1).sum()), "overflow": int((r.status == 2).sum())},
                  "cpu_oracle": {"searches": len(sample), "wall_s": t_cpu, "expanded_nodes_per_sec": exp_cpu / t_cpu, "cores": 1,
                                 "note": "oracle/astar_oracle.py (pure Python + numpy, as the reference planner is); identical sequences and expansion counts asserted on the sample"}}))
