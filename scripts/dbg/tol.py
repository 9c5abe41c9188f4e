import sys, os
sys.path.insert(0, os.getcwd()); sys.path.insert(0, 'tests'); sys.path.insert(0, 'oracle')
import numpy as np
from conftest import golden
import test_gpu_parity_step as t
from mpc_for_av_at_intersection_b200 import MpcConfig
for name in ("random_T20.npz", "closed_loop_T20.npz"):
    g = golden(name)
    sel = np.arange(len(g["state"]))
    s = g["solved"].astype(bool)
    for tm, tp, td in ((1e-12, 1e-10, 1e-9), (1e-11, 1e-10, 1e-9), (1e-10, 1e-9, 1e-8), (1e-9, 1e-8, 1e-7), (1e-8, 1e-8, 1e-7)):
        cfg = MpcConfig.from_json(T=20, ipm_tol_mu=tm, ipm_tol_p=tp, ipm_tol_d=td)
        out = t._run_session(cfg, g, sel)
        e = np.maximum(np.abs(out["di_ai"][:, 0] - g["di"]), np.abs(out["di_ai"][:, 1] - g["ai"]))
        print(name, tm, f"max|du0|={e[s].max():.2e} p99={np.quantile(e[s], .99):.2e} iters={out['iters'][s].mean():.2f} status1={(out['status']==1).sum()}")
