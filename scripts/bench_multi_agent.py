"""BASELINE.json configs[2]: 1024 scenarios x 8 interacting vehicles, T = 20, closed loop with inter-vehicle coupling
(coupling_kernel + mpc_step_kernel per control step, everything resident on the device).
usage: python scripts/bench_multi_agent.py [scenarios] [steps]  ->  one JSON line"""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mpc_for_av_at_intersection_b200 import CourseSet, MpcConfig, MpcSession, load_intersection_routes
from mpc_for_av_at_intersection_b200.scenarios import make_multi_agent_workload
S = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
K = int(sys.argv[2]) if len(sys.argv) > 2 else 150
V, W = 8, 5
cfg = MpcConfig.from_json(T=20)
cs = CourseSet(load_intersection_routes())
wl = make_multi_agent_workload(cs, S, V, seed=1236)
B = S * V
ses = MpcSession(cfg, cs, B).begin(wl["route"], target_ind=wl["target_ind"])
ses.store("state", wl["state"])
st = torch.cuda.Stream()
ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
active, cuts = [], []
with torch.cuda.stream(st):
    for k in range(W):
        ses.multi_agent_dev(1, V, 0, st.cuda_stream)
    st.synchronize()
    l0 = ses.launches
    t0 = time.perf_counter()
    for k in range(K):
        ev[k][0].record(st)
        ses.multi_agent_dev(1, V, 0, st.cuda_stream)
        ev[k][1].record(st)
        if k % 10 == 0:
            st.synchronize()
            done = ses.fetch("done"); nc = ses.fetch("ncourse")
            active.append(int((done == 0).sum())); cuts.append(int((nc < cs.length[wl["route"]]).sum()))
    st.synchronize()
    t1 = time.perf_counter()
ms = np.array([a.elapsed_time(b) for a, b in ev])
done = ses.fetch("done")
# a vehicle that reached its goal is skipped by both kernels: count the solves that actually ran
status_runs = None
print(json.dumps({"metric": "vehicle_steps_per_sec", "workload": f"BASELINE.json configs[2]: {S} scenarios x {V} vehicles, T=20, {K} control steps after {W} warm-up steps, coupling per SURVEY 8a A9",
                  "value": float(np.sum(active and np.interp(np.arange(K), np.arange(0, K, 10), active)) / (ms.sum() / 1e3)),
                  "unit": "active vehicle-steps/s", "ms_per_step_mean": float(ms.mean()), "ms_per_step_p50": float(np.median(ms)),
                  "ms_first_steps": ms[:5].round(3).tolist(), "active_vehicles_every_10_steps": active, "cut_paths_every_10_steps": cuts,
                  "reached_goal": int((done == 1).sum()), "dead": int((done == 2).sum()), "launches": ses.launches - l0,
                  "wall_s": t1 - t0}))
