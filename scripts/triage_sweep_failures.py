"""Development tool: re-run one shard of the closed-loop sweep (bench.py --workload sweep) step by step and replay every solve the
device reports as FAILED through the C++ port on identical inputs -- 'kernel fails, port solves' must be 0.  Dumps the
offending problems to gpurun_out/ for scripts/trace_failures.py / scripts/replay_twin.py.
usage (GPU box): python scripts/triage_sweep_failures.py T seed scenarios [steps]     (TRIAGE_L=<wheelbase>, TRIAGE_GOAL_TEST=0 for the
config-5 generator: scripts/sweep_config5.py uses seed 1238, no goal test)"""
import os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO); sys.path.insert(0, os.path.join(REPO, "oracle"))
import mpc_port
from mpc_for_av_at_intersection_b200 import CourseSet, MpcConfig, MpcSession, load_intersection_routes
from mpc_for_av_at_intersection_b200.scenarios import make_workload

T, seed, n = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 80
cs = CourseSet(load_intersection_routes())
cfg = MpcConfig.from_json(T=T)
cfgvec = mpc_port.pack_config(cfg)
ses = MpcSession(cfg, cs, n)
wl = make_workload(cs, n, seed=seed)
Lw = float(os.environ.get("TRIAGE_L", cfg.wheelbase))
goal_test = os.environ.get("TRIAGE_GOAL_TEST", "1") != "0"
ses.begin(wl["route"], target_ind=wl["target_ind"], L=np.full(n, Lw))
ses.store("state", wl["state"])
PRE = ("state", "oa", "odelta", "target_ind", "warm", "di_ai", "ncourse", "route", "done")
total = kernel_only = raised = 0
os.makedirs("gpurun_out", exist_ok=True)
for k in range(steps):
    pre = {name: ses.fetch(name) for name in PRE}
    ses.closed_loop_dev(1, 0, 0, goal_test=goal_test)
    status, done = ses.fetch("status"), ses.fetch("done")
    ran = (pre["done"] == 0) & (done != 1)
    near = np.flatnonzero(ran & (status == 3))          # the reference's nearest-index Exception: not a failed solve
    if len(near):
        c3 = lambda a: np.ascontiguousarray(a[near])
        t3, w3, a3, d3, u3 = c3(pre["target_ind"]), c3(pre["warm"]), c3(pre["oa"]), c3(pre["odelta"]), c3(pre["di_ai"])
        o3 = mpc_port.step_batch(cfgvec, cs.packed, cs.dl, c3(pre["state"]), c3(pre["route"]), c3(pre["ncourse"]), np.full(len(near), Lw), t3, w3, a3, d3, u3)
        print(f"step {k}: {len(near)} nearest-index Exceptions on the device (problems {near.tolist()}), port statuses {o3['status'].tolist()}", flush=True)
        raised += len(near)
    bad = np.flatnonzero(ran & (status == 2))
    total += len(bad)
    if len(bad):
        c = lambda a: np.ascontiguousarray(a[bad])
        tgt, warm, oa, od, di_ai = c(pre["target_ind"]), c(pre["warm"]), c(pre["oa"]), c(pre["odelta"]), c(pre["di_ai"])
        out = mpc_port.step_batch(cfgvec, cs.packed, cs.dl, c(pre["state"]), c(pre["route"]), c(pre["ncourse"]), np.full(len(bad), Lw),
                                  tgt, warm, oa, od, di_ai)
        wrong = out["status"] <= 1
        kernel_only += int(wrong.sum())
        print(f"step {k}: {len(bad)} failed on the device, port statuses {out['status'].tolist()}, port iterations {out['iters'].tolist()}", flush=True)
        if wrong.any():
            np.savez(f"gpurun_out/sweep_kernel_only_T{T}_s{seed}_step{k}.npz", **{name: pre[name][bad[wrong]] for name in PRE if name != "done"},
                     port_di_ai=di_ai[wrong], gpu_iters=ses.fetch("iters")[bad[wrong]])
print(f"T={T} seed={seed} scenarios={n} steps={steps}: {total} failed on the device, {kernel_only} of them solved by the port; {raised} nearest-index Exceptions")
