"""Bitwise comparison of the step kernel across tuning variants and repeated runs (needs a MPCB_TUNE_VARIANTS build)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mpc_for_av_at_intersection_b200 import CourseSet, MpcConfig, MpcSession, load_intersection_routes
from mpc_for_av_at_intersection_b200.scenarios import make_workload
B = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 6
cfg = MpcConfig.from_json(T=20)
cs = CourseSet(load_intersection_routes())
wl = make_workload(cs, B, seed=1235)
def run(variant):
    os.environ["MPCB_VARIANT"] = str(variant)
    ses = MpcSession(cfg, cs, B).begin(wl["route"], target_ind=wl["target_ind"])
    ses.store("state", wl["state"])
    out = []
    for k in range(steps):
        ses.closed_loop_dev(1, 0, 0, goal_test=False)
        out.append((ses.fetch("di_ai").copy(), ses.fetch("status").copy(), ses.fetch("iters").copy(), ses.fetch("state").copy()))
    ses.close()
    return out
ref = run(0)
for v in [0] + [int(x) for x in os.environ.get("VARIANTS", "1 2 3").split()]:
    o = run(v)
    for k in range(steps):
        nd = int((o[k][0] != ref[k][0]).any(axis=1).sum()); ns = int((o[k][1] != ref[k][1]).sum()); ni = int((o[k][2] != ref[k][2]).sum())
        mx = float(np.abs(o[k][0] - ref[k][0]).max())
        print(f"variant {v} step {k}: di_ai rows differing {nd} (max abs {mx:.3e}) status {ns} iters {ni} fails {int((o[k][1] >= 2).sum())}", flush=True)
