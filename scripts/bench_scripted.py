"""SURVEY 8f N1 / N3 measurements (one JSON line each):
  scripted  -- the reference's whole loop body with scripted cross traffic on the device (coupling_kernel +
               mpc_step_kernel per control step), launched eagerly and replayed as ONE CUDA graph;
  variants  -- the variant kernel (per-problem weight rows + speed reference) against the default kernel on the
               bench workload of BASELINE.json configs[1].
usage: python scripts/bench_scripted.py [egos] [steps]"""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mpc_for_av_at_intersection_b200 import CourseSet, MpcConfig, MpcSession, load_intersection_routes
from mpc_for_av_at_intersection_b200.scenarios import make_scripted_workload, make_workload

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
K = int(sys.argv[2]) if len(sys.argv) > 2 else 100
cfg = MpcConfig.from_json(T=20)
cs = CourseSet(load_intersection_routes())
st = torch.cuda.Stream()


def timed(fn, n):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    st.synchronize()
    a.record(st); fn(); b.record(st)
    st.synchronize()
    return a.elapsed_time(b) / n


# ---- scripted loop --------------------------------------------------------------------------------------------
wl = make_scripted_workload(cs, B)
ses = MpcSession(cfg, cs, B)


def reset():
    ses.begin(wl["route"], target_ind=wl["target_ind"])
    ses.store("state", wl["state"])
    ses.set_scripted_obstacles(wl["scripts"])


reset()
with torch.cuda.stream(st):
    ses.scripted_loop_dev(3, 0, 0, st.cuda_stream)            # warm-up (function attributes, caches)
    reset()
    l0 = ses.launches
    ms_eager = timed(lambda: ses.scripted_loop_dev(K, 0, 0, st.cuda_stream), K)
    launches = ses.launches - l0
    done_e, nc_e, state_e = ses.fetch("done"), ses.fetch("ncourse"), ses.fetch("state")
    reset()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=st):
        ses.scripted_loop_dev(K, 0, 0, torch.cuda.current_stream().cuda_stream)
    ms_graph = timed(g.replay, K)
    done_g, nc_g, state_g = ses.fetch("done"), ses.fetch("ncourse"), ses.fetch("state")
assert np.array_equal(done_e, done_g) and np.array_equal(nc_e, nc_g) and np.array_equal(state_e, state_g), "graph replay must reproduce the eager loop"
print(json.dumps({"metric": "scripted_loop_ms_per_control_step", "workload": f"{B} egos on the 12 routes, 0-3 scripted crossing cars each (make_scripted_workload), T=20, {K} control steps from the route heads",
                  "eager_ms": ms_eager, "graph_ms": ms_graph, "launches_eager": launches, "graph_launches": 1,
                  "vehicle_steps_per_s_graph": B / (ms_graph * 1e-3), "reached_goal": int((done_g == 1).sum()),
                  "cut_paths_at_end": int((nc_g < cs.length[wl["route"]]).sum()), "graph_equals_eager": True}))
ses.close()

# ---- variant kernel vs default kernel on the bench workload ---------------------------------------------------------
Bv = 4096
wl = make_workload(cs, Bv)
res = {}
for name in ("default", "weights", "weights+speed_ref"):
    c = MpcConfig.from_json(T=20, ref_speed=25 / 3.6 if "speed" in name else 0.0, Q_v_yaw=[20.0 if "speed" in name else 0.0, 0.5])
    s = MpcSession(c, cs, Bv).begin(wl["route"], target_ind=wl["target_ind"])
    s.store("state", wl["state"])
    if name != "default":
        s.set_weights(np.tile(c.weight_row(), (Bv, 1)))
    with torch.cuda.stream(st):
        s.closed_loop_dev(5, 0, st.cuda_stream, goal_test=False)
        res[name] = timed(lambda: s.closed_loop_dev(40, 0, st.cuda_stream, goal_test=False), 40)
    res[name + "_iters"] = float(s.fetch("iters").mean())
    s.close()
print(json.dumps({"metric": "variant_kernel_ms_per_step", "workload": f"BASELINE.json configs[1] workload, B={Bv}, T=20, 40 closed-loop steps", **res}))
