"""Replay dumped problems (gpurun_out/*.npz from tests/test_gpu_bench_workloads.py) through the numpy twin of the kernel's
interior point (oracle/condensed_model.py) and the numpy oracle, on a CPU-only box.
usage: python scripts/replay_twin.py file.npz T [index] [-v]"""
import os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO); sys.path.insert(0, os.path.join(REPO, "oracle")); sys.path.insert(0, os.path.join(REPO, "tests"))
import mpc_oracle as mo
import condensed_model as cm
from conftest import load_routes


def build(d, i, T, L=float(os.environ.get("REPLAY_L", "2.86"))):
    routes = load_routes(True)
    full = routes[int(d["route"][i])][: int(d["ncourse"][i])]
    x, y, yaw, v = d["state"][i]
    cfg = mo.Cfg.from_dict({"T": T})
    dl = float(np.linalg.norm(routes[int(d["route"][i])][0, :2] - routes[int(d["route"][i])][1, :2]))
    oa = d["oa"][i] if d["warm"][i] else np.zeros(T)
    od = d["odelta"][i] if d["warm"][i] else np.zeros(T)
    xref, tind, idx, re = mo.calc_ref_trajectory(x, y, v, full[:, 0], full[:, 1], full[:, 2], dl, 0.2, int(d["target_ind"][i]), T, None, None)
    xbar = mo.predict_motion([x, y, v, yaw], list(oa), list(od), L, 0.2, T)
    c = cm.Condensed(xref, xbar, [x, y, v, yaw], re, cfg, L, 0.2)
    return c, oa, od, (xref, xbar, [x, y, v, yaw], re, cfg)


if __name__ == "__main__":
    d = np.load(sys.argv[1]); T = int(sys.argv[2])
    ids = [int(sys.argv[3])] if len(sys.argv) > 3 and sys.argv[3].isdigit() else range(len(d["state"]))
    for i in ids:
        c, oa, od, raw = build(d, i, T)
        r = c.solve(oa, od, max_iter=60, verbose="-v" in sys.argv)
        sol = mo.linear_mpc_control(*raw, float(os.environ.get("REPLAY_L", "2.86")), 0.2)
        print(f"problem {i}: twin status {r['status']} iters {r['iters']} u0=({r['od'][0]:.9f},{r['oa'][0]:.9f}); oracle:",
              None if sol is None else (sol["status"], sol["iters"], float(sol["od"][0]), float(sol["oa"][0])),
              "gpu iters", d["gpu_iters"][i] if "gpu_iters" in d.files else None,
              "port u0", d["port_di_ai"][i] if "port_di_ai" in d.files else None)
