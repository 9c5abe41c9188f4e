"""VERDICT r01 item 2, last bullet: "FP32 LDL' of H + G'WG + FP64 residual refinement -- report it even if slower".
Answered on the numpy twin of the kernel's interior point (oracle/condensed_model.py), CPU only: closed loops of the bench
generator (config 2, T = 20), every solve once with the float64 factorisation and once with a float32 Cholesky of the
Jacobi-scaled matrix + k refinement steps against the float64 matrix.  usage: python scripts/fp32_factor_experiment.py [n] [steps]"""
import os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO); sys.path.insert(0, os.path.join(REPO, "oracle")); sys.path.insert(0, os.path.join(REPO, "tests"))
import mpc_oracle as mo
import condensed_model as cm
from conftest import load_routes
from mpc_for_av_at_intersection_b200.courses import CourseSet
from mpc_for_av_at_intersection_b200.scenarios import make_workload

n = int(sys.argv[1]) if len(sys.argv) > 1 else 48
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 15
T, L, dt = 20, 2.86, 0.2
routes = load_routes(True)
cs = CourseSet(load_routes(False))
wl = make_workload(cs, n, seed=1235)
cfg = mo.Cfg.from_dict({"T": T})
variants = [("f64", None), ("f32 + 0 refinement", ("f32", 0)), ("f32 + 1", ("f32", 1)), ("f32 + 2", ("f32", 2)), ("f32 + 4", ("f32", 4))]
stat = {name: dict(it=[], fail=0, err=[], cond=[]) for name, _ in variants}
for b in range(n):
    r = routes[wl["route"][b]]
    dl = float(np.linalg.norm(r[0, :2] - r[1, :2]))
    x, y, yaw, v = wl["state"][b]
    tind = int(wl["target_ind"][b])
    oa = od = None
    for k in range(steps):
        xref, tind, idx, re = mo.calc_ref_trajectory(x, y, v, r[:, 0], r[:, 1], r[:, 2], dl, dt, tind, T, None, None)
        a0 = np.zeros(T) if oa is None else oa
        d0 = np.zeros(T) if od is None else od
        xbar = mo.predict_motion([x, y, v, yaw], list(a0), list(d0), L, dt, T)
        c = cm.Condensed(xref, xbar, [x, y, v, yaw], re, cfg, L, dt)
        ref = None
        for name, fac in variants:
            res = c.solve(a0, d0, max_iter=100, factor=fac)
            if k >= 5:
                st = stat[name]
                st["it"].append(res["iters"])
                if res["status"] != 0:
                    st["fail"] += 1
                elif ref is not None:
                    st["err"].append(max(abs(res["oa"][0] - ref["oa"][0]), abs(res["od"][0] - ref["od"][0])))
            if fac is None:
                ref = res
        oa, od = ref["oa"], ref["od"]
        x, y, yaw, v = mo.plant_step(x, y, yaw, v, oa[0], od[0], L, dt)
for name, _ in variants:
    st = stat[name]
    e = np.array(st["err"]) if st["err"] else np.array([0.0])
    print(f"{name:20s} solves {len(st['it'])}  iterations/solve {np.mean(st['it']):.2f}  not converged {st['fail']}  "
          f"|du0| vs f64: p50 {np.percentile(e, 50):.1e} p99 {np.percentile(e, 99):.1e} max {e.max():.1e}")
