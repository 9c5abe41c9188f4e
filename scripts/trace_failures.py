"""Development tool: replay dumped problems (gpurun_out/*.npz of tests/test_gpu_bench_workloads.py) through a trace build of
the library (-DMPCB_TRACE) and print the interior-point log of every problem.
usage (GPU box): MPCB_EXTRA_NVCC_FLAGS=-DMPCB_TRACE python scripts/trace_failures.py T file.npz [file.npz ...]"""
import ctypes as C, glob, os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import torch
from mpc_for_av_at_intersection_b200 import CourseSet, MpcConfig, _lib, load_intersection_routes
from mpc_for_av_at_intersection_b200.batched import BatchedMPC

T = int(sys.argv[1])
files = sum((sorted(glob.glob(a)) for a in sys.argv[2:]), [])
os.utime(os.path.join(_lib.CSRC, "mpcb_api.cu"))
_lib.build()
d = [np.load(f) for f in files]
cat = lambda k, dt: np.ascontiguousarray(np.concatenate([x[k] for x in d]).astype(dt))
state, route, nc = cat("state", np.float64), cat("route", np.int32), cat("ncourse", np.int32)
tgt, warm, oa, od, di_ai = cat("target_ind", np.int32), cat("warm", np.int32), cat("oa", np.float64), cat("odelta", np.float64), cat("di_ai", np.float64)
B = len(state)
cfg = MpcConfig.from_json(T=T)
if len(sys.argv) and os.environ.get("IPM_MAX_ITER"):
    cfg.ipm_max_iter = int(os.environ["IPM_MAX_ITER"])
eng = BatchedMPC(cfg, CourseSet(load_intersection_routes()), B)
trace = torch.zeros(B, 128, 8, dtype=torch.float64, device="cuda")
eng._lib.mpcb_debug_set_trace.argtypes = [C.c_void_p]
eng._lib.mpcb_debug_set_trace(C.c_void_p(trace.data_ptr()))
Lw = np.full(B, float(os.environ.get("REPLAY_L", "2.86")))          # wheelbase of every replayed problem
out = eng.step_host(state, route, nc, tgt, warm, oa, od, di_ai, L=Lw, want_traj=False)
tr = trace.cpu().numpy()
for b in range(B):
    print(f"== problem {b}: status {out['status'][b]} iters {out['iters'][b]} u0 = {di_ai[b]}")
    for it in range(out["iters"][b] + 1):
        r = tr[b, it]
        print(f"  {int(r[0]):3d} res_d {r[1]:.3e} res_p {r[2]:.3e} mu {r[3]:.3e} al {r[4]:.4f} verified {int(r[5])} bad {int(r[6])} gscale {r[7]:.2e}")
