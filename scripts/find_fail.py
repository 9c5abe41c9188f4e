"""Find solves that end with status >= 2 in the bench workload and save their inputs (gpurun_out/fail.npz)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mpc_for_av_at_intersection_b200 import CourseSet, MpcConfig, MpcSession, load_intersection_routes
from mpc_for_av_at_intersection_b200.scenarios import make_workload
B = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 30
cfg = MpcConfig.from_json(T=20)
cs = CourseSet(load_intersection_routes())
wl = make_workload(cs, B, seed=1235)
ses = MpcSession(cfg, cs, B).begin(wl["route"], target_ind=wl["target_ind"])
ses.store("state", wl["state"])
found = []
for k in range(steps):
    pre = {n: ses.fetch(n) for n in ("state", "oa", "odelta", "target_ind", "warm", "di_ai", "ncourse", "route")}
    ses.closed_loop_dev(1, 0, 0, goal_test=False)
    st, it = ses.fetch("status"), ses.fetch("iters")
    bad = np.flatnonzero(st >= 2)
    print(f"step {k}: iters mean {it.mean():.2f} max {it.max()} status>=2: {len(bad)}", flush=True)
    for b in bad:
        found.append({n: pre[n][b] for n in pre} | {"step": k, "b": int(b), "status": int(st[b]), "iters": int(it[b])})
if found:
    np.savez("gpurun_out/fail.npz", **{n: np.array([f[n] for f in found]) for n in found[0]})
    print("saved", len(found), "failures")
