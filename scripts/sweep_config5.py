"""BASELINE.json configs[4]: horizon x batch x wheelbase sweep of the fused step (latency at small B, throughput at large B),
ONE process: T in {10, 13, 20, 40, 80} x B in {1, 2, 4, ..., 65536} x L in {2.86 (BicycleModelDimensions), 4.0 (PriusDimensions
at scale 1: main/lib/car_dimensions.py:82-107)}.  Per cell: the config-2 generator's closed loops from the cold start, 5 untimed
lead-in steps, then timed control steps 5.. (20, fewer for the multi-second cells), one fused launch per step, CUDA events per
step, L2 flushed between steps.  Reports mean / p50 / p99 per-step latency and solves/s.  Also the B = 1 latency a scenario
script sees through compat.mpc.MPC.step (config 1's view: host State in, (di, ai) and the object's attributes out).

usage: python scripts/sweep_config5.py out.json [--quick]"""
import json
import os
import sys
import time

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import torch  # noqa: E402
from mpc_for_av_at_intersection_b200 import CourseSet, MpcConfig, MpcSession, load_intersection_routes  # noqa: E402
from mpc_for_av_at_intersection_b200.scenarios import make_workload  # noqa: E402

out_path = sys.argv[1] if len(sys.argv) > 1 and not sys.argv[1].startswith("-") else "gpurun_out/config5_sweep.json"
quick = "--quick" in sys.argv
LEAD = 5
RATE = {10: 9e6, 13: 8e6, 20: 6e6, 40: 7e5, 80: 1e5}          # rough solves/s, only to bound the time per cell
dev = torch.device("cuda", 0)
cs = CourseSet(load_intersection_routes())
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
stream = torch.cuda.Stream(device=dev)
rows = []
Bs = [1 << k for k in range(0, 17, 4 if quick else 1)]
for T in (10, 13, 20, 40, 80):
    cfg = MpcConfig.from_json(T=T)
    ses = MpcSession(cfg, cs, max(Bs))
    for L in (2.86, 4.0):
        for B in Bs:
            steps = 20 if B * 25 / RATE[T] < 2.0 else 6
            wl = make_workload(cs, B, seed=1234 + 4)
            ses.begin(wl["route"], target_ind=wl["target_ind"], L=np.full(B, L))
            ses.store("state", wl["state"])
            iters_v = torch.as_tensor(ses.device_view("iters"), device=dev)
            status_v = torch.as_tensor(ses.device_view("status"), device=dev)
            ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
            it_acc = torch.zeros((), dtype=torch.float64, device=dev)
            bad = torch.zeros((), dtype=torch.int64, device=dev)
            exc = torch.zeros((), dtype=torch.int64, device=dev)     # status 3: the reference's nearest-index Exception, not a failed solve
            with torch.cuda.stream(stream):
                for _ in range(LEAD):
                    ses.closed_loop_dev(1, 0, stream.cuda_stream, goal_test=False)
                for k in range(steps):
                    flush.zero_()
                    ev[k][0].record(stream)
                    ses.closed_loop_dev(1, 0, stream.cuda_stream, goal_test=False)
                    ev[k][1].record(stream)
                    it_acc += iters_v.double().mean()
                    bad += (status_v == 2).sum()
                    exc += (status_v == 3).sum()
            torch.cuda.synchronize(dev)
            ms = np.array([a.elapsed_time(b) for a, b in ev])
            rows.append(dict(T=T, B=B, L=L, steps=steps, ms_mean=float(ms.mean()), ms_p50=float(np.percentile(ms, 50)),
                             ms_p99=float(np.percentile(ms, 99)), solves_per_s=float(B * steps / ms.sum() * 1e3),
                             mean_ipm_iters=float(it_acc.item()) / steps, failed_solves=int(bad.item()),
                             nearest_index_exceptions=int(exc.item())))
            print(rows[-1], flush=True)
    ses.close()

# ---- config 1's view: one vehicle through the reference-shaped facade (default JSON: T = 13) ----------------------------
from mpc_for_av_at_intersection_b200.compat import mpc as cmpc  # noqa: E402
from mpc_for_av_at_intersection_b200.compat.car_dimensions import BicycleModelDimensions  # noqa: E402
from mpc_for_av_at_intersection_b200.compat.simulation import HistorySimulation, State  # noqa: E402
full = load_intersection_routes()[3 * 3 + 1] if len(load_intersection_routes()) >= 12 else load_intersection_routes()[0]
car = BicycleModelDimensions(skip_back_circle_collision_checking=False)
mpc = cmpc.MPC(cx=full[:, 0], cy=full[:, 1], cyaw=full[:, 2], dl=np.linalg.norm(full[0, :2] - full[1, :2]), dt=0.2, car_dimensions=car)
state = State(x=full[0, 0], y=full[0, 1], yaw=full[0, 2], v=0.0)
sim = HistorySimulation(car_dimensions=car, sample_time=0.2, initial_state=state)
lat = []
for k in range(60):
    if mpc.is_goal(state):
        break
    t0 = time.perf_counter()
    delta, acc = mpc.step(state)
    lat.append((time.perf_counter() - t0) * 1e3)
    state = sim.step(a=acc, delta=delta, xref_deviation=mpc.get_current_xref_deviation())
lat = np.array(lat[5:])
facade = dict(T=int(cmpc.T), steps=len(lat), ms_mean=float(lat.mean()), ms_p50=float(np.percentile(lat, 50)), ms_p99=float(np.percentile(lat, 99)),
              what="compat.mpc.MPC.step wall clock: host State in, one mpcb_session_step_all_host call (H2D, kernel, D2H of di/ai, status, "
                   "target_ind, oa, odelta, ox..oyaw, xref, one synchronisation), attributes set")
print("facade", facade, flush=True)
json.dump(dict(config="BASELINE.json configs[4]: cfg-2 generator (seed 1238), closed loop, fused step + plant, timed control steps 5.., "
                      "L2 flushed between steps; L = wheelbase of every vehicle (2.86 bicycle model, 4.0 Prius at scale 1)",
               rows=rows, facade_b1=facade), open(out_path, "w"), indent=1)
