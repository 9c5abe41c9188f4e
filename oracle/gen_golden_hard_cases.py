"""TEST INFRASTRUCTURE -- assemble tests/golden/hard_cases.npz from the problems that tests/test_gpu_bench_workloads.py dumped
when the round-1 kernels failed them (gpurun_in/*.npz: one config-2 step with v0 = MAX_SPEED whose carried dual residual
drifted, one T = 40 and ten T = 80 steps of vehicles far off their path: cost gradients of 1e6..1e7, 60+ tiny interior-point
steps from the old lam = 1 start, or a pivot breakdown at mu ~ 1e-12).  The expected answers are those of the independent
C++ port (oracle/mpc_port.cpp), which solves all of them.  usage: python oracle/gen_golden_hard_cases.py gpurun_in"""
import glob, os, sys
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE); sys.path.insert(0, os.path.dirname(HERE))
import mpc_port
from mpc_for_av_at_intersection_b200 import CourseSet, MpcConfig, load_intersection_routes

src = sys.argv[1] if len(sys.argv) > 1 else "gpurun_in"
cs = CourseSet(load_intersection_routes(), smooth="numpy")
rows = []
for f in sorted(glob.glob(os.path.join(src, "*.npz"))):
    T = 20 if "cfg2" in f else int(os.path.basename(f).split("_T")[1].split("_")[0])
    d = np.load(f)
    cfg = MpcConfig.from_json(T=T)
    n = len(d["state"])
    state, route, nc = d["state"].copy(), d["route"].astype(np.int32), d["ncourse"].astype(np.int32)
    tgt, warm, oa, od, di_ai = d["target_ind"].astype(np.int32), d["warm"].astype(np.int32), d["oa"].copy(), d["odelta"].copy(), d["di_ai"].copy()
    out = mpc_port.step_batch(mpc_port.pack_config(cfg), cs.packed, cs.dl, state, route, nc, np.full(n, 2.86), tgt, warm, oa, od, di_ai, want_traj=False)
    for i in range(n):
        pad = lambda v: np.concatenate([v, np.zeros(80 - T)])
        rows.append(dict(T=T, state=d["state"][i], route=d["route"][i], ncourse=d["ncourse"][i], target_ind=d["target_ind"][i], warm=d["warm"][i],
                         oa=pad(d["oa"][i]), odelta=pad(d["odelta"][i]), di_ai=d["di_ai"][i], port_di_ai=di_ai[i], port_status=out["status"][i],
                         port_iters=out["iters"][i], port_target=tgt[i]))
        print(os.path.basename(f), i, "T", T, "port status", out["status"][i], "iters", out["iters"][i], "u0", di_ai[i])
np.savez_compressed(os.path.join(os.path.dirname(HERE), "tests", "golden", "hard_cases.npz"), **{k: np.array([r[k] for r in rows]) for k in rows[0]})
print("saved", len(rows))
