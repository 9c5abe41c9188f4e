"""TEST INFRASTRUCTURE -- assemble tests/golden/hard_cases.npz from the problems that tests/test_gpu_bench_workloads.py dumped
when the round-1 kernels failed them (gpurun_in/*.npz: one config-2 step with v0 = MAX_SPEED whose carried dual residual
drifted, one T = 40 and ten T = 80 steps of vehicles far off their path: cost gradients of 1e6..1e7, 60+ tiny interior-point
steps from the old lam = 1 start, or a pivot breakdown at mu ~ 1e-12).  The expected answers are those of the independent
C++ port (oracle/mpc_port.cpp), which solves all of them.  usage: python oracle/gen_golden_hard_cases.py gpurun_in
Later finds are appended: python oracle/gen_golden_hard_cases.py --append dump_T80_x.npz [more dumps]   (round 2: one T = 80 step of
the 32 768-scenario sweep with seed 1, scripts/triage_sweep_failures.py -- the port needs 42 iterations and reports its
inaccurate class, the numpy oracle and the numpy twin give up; the kernels ended FAILED at a gap of 2e-7 before the gap
was judged relative to the cost scale)."""
import glob, os, sys
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE); sys.path.insert(0, os.path.dirname(HERE))
import mpc_port
from mpc_for_av_at_intersection_b200 import CourseSet, MpcConfig, load_intersection_routes

GOLD = os.path.join(os.path.dirname(HERE), "tests", "golden", "hard_cases.npz")
append = len(sys.argv) > 1 and sys.argv[1] == "--append"
src = sys.argv[1] if len(sys.argv) > 1 else "gpurun_in"
cs = CourseSet(load_intersection_routes(), smooth="numpy")
rows = []
if append:
    g = np.load(GOLD)
    rows = [{k: g[k][i] for k in g.files} for i in range(len(g["T"]))]
for f in (sys.argv[2:] if append else sorted(glob.glob(os.path.join(src, "*.npz")))):
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
np.savez_compressed(GOLD, **{k: np.array([r[k] for r in rows]) for k in rows[0]})
print("saved", len(rows))
