"""TEST / MEASUREMENT INFRASTRUCTURE -- time the CPU restatement of the reference's MPC loop.

Used only by bench.py (`cpu_baseline` leg and `--impl reference`).  The unmodified reference
(/root/reference, Python + cvxpy -> ECOS) cannot travel to the GPU box and its solver stack is absent, so
the CPU arm is the oracle port (oracle/mpc_oracle.py: the reference's own non-condensed QP, float64 IPM):
`kind = "port"`.  One process per host core (fork), BLAS pinned to one thread per process, each worker runs
closed loops `mpc.step -> Simulation.step` (main/scenarios/mpc_intersection.py:142,159) on its share of the
sample."""
import multiprocessing as mp
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)

_G = {}


def _worker(args):
    ids, steps = args
    try:
        from threadpoolctl import threadpool_limits
        threadpool_limits(1)
    except Exception:
        pass
    import mpc_oracle as mo
    packed, length, dl, route, state, target, T, L, dt = (_G[k] for k in
                                                           ("packed", "length", "dl", "route", "state", "target", "T", "L", "dt"))
    cfg = mo.Cfg.from_dict({"T": T})
    n_solves = 0
    out = []
    t0 = time.perf_counter()
    for b in ids:
        r = int(route[b])
        n = int(length[r])
        om = mo.OracleMPC(packed[r, 0, :n].copy(), packed[r, 1, :n].copy(), packed[r, 2, :n].copy(),
                          float(dl[r]), L, dt, cfg)
        om.target_ind = int(target[b])
        st = list(state[b])
        for _ in range(steps):
            try:
                di, ai = om.step(st)
            except Exception:
                break
            n_solves += 1
            st = list(mo.plant_step(st[0], st[1], st[2], st[3], ai, di, L, dt))
        out.append((b, om.di, om.ai))
    return n_solves, time.perf_counter() - t0, out


def run(courses, workload, ids, steps, T, L, dt, cores=None):
    """Closed loops of `steps` control steps for problems `ids` of `workload` over `cores` processes.
    Returns dict(solves, seconds (wall), cores, per_core_busy_s, controls)."""
    cores = int(cores or os.cpu_count() or 1)
    _G.update(packed=courses.packed, length=courses.length, dl=courses.dl, route=workload["route"],
              state=workload["state"], target=workload["target_ind"], T=T, L=L, dt=dt)
    ids = list(ids)
    cores = max(1, min(cores, len(ids)))
    chunks = [ids[i::cores] for i in range(cores)]
    t0 = time.perf_counter()
    if cores == 1:
        res = [_worker((chunks[0], steps))]
    else:
        ctx = mp.get_context("fork")
        with ctx.Pool(cores) as pool:
            res = pool.map(_worker, [(c, steps) for c in chunks])
    wall = time.perf_counter() - t0
    solves = sum(r[0] for r in res)
    controls = sorted(sum((r[2] for r in res), []))
    return dict(solves=solves, seconds=wall, cores=cores, per_core_busy_s=[r[1] for r in res], controls=controls)


def run_cpp(courses_smoothed, workload, n, steps, cfg, cores=None, min_seconds=0.0, max_steps=2000, lead_in=5, phase_len=20):
    """Same closed loops through the plain C++ port (oracle/mpc_port.cpp), `cores` std::threads inside one call per
    control step; the plant step between calls is the vectorised host twin of Simulation.step (not timed: in the
    reference the scenario script owns the simulation).  Same step mix as the GPU arm of bench.py: every cycle starts
    the closed loops from the workload's cold start, runs `lead_in` untimed control steps and then times up to
    `phase_len` steps (control steps lead_in .. lead_in + phase_len - 1).  Only `step_batch` is inside the clock.
    `courses_smoothed`: CourseSet with smoothed yaws (the port, like MPC.__init__, works on smoothed yaws).
    Runs at least `steps` timed control steps and keeps going until `min_seconds` of timed work have passed.
    Returns dict(solves, seconds (timed), cores, failed, steps, wall)."""
    import mpc_port
    from mpc_for_av_at_intersection_b200.scenarios import host_plant_step
    cores = int(cores or os.cpu_count() or 1)
    T = int(cfg.T)
    cfgvec = mpc_port.pack_config(cfg)
    mpc_port.lib()
    route = np.ascontiguousarray(workload["route"][:n]).astype(np.int32)
    ncourse = courses_smoothed.length[route].astype(np.int32)
    L = np.full(n, float(cfg.wheelbase))
    solves = failed = 0
    timed = 0.0
    t_wall = time.perf_counter()
    k = 0
    while k < steps or (timed < min_seconds and k < max_steps):
        state = np.ascontiguousarray(workload["state"][:n]).copy()           # the same cold start every cycle
        target = np.ascontiguousarray(workload["target_ind"][:n]).astype(np.int32)
        warm = np.zeros(n, np.int32)
        oa = np.zeros((n, T)); od = np.zeros((n, T)); di_ai = np.zeros((n, 2))
        for j in range(lead_in + phase_len):
            is_timed = j >= lead_in
            if is_timed and not (k < steps or (timed < min_seconds and k < max_steps)):
                break
            t0 = time.perf_counter()
            out = mpc_port.step_batch(cfgvec, courses_smoothed.packed, courses_smoothed.dl, state, route, ncourse, L, target, warm,
                                      oa, od, di_ai, want_traj=False, nthreads=cores)
            if is_timed:
                timed += time.perf_counter() - t0
                k += 1
                solves += n
                failed += int((out["status"] >= 2).sum())
            state = host_plant_step(state, di_ai[:, 0], di_ai[:, 1], float(cfg.wheelbase), float(cfg.dt))
    return dict(solves=solves, seconds=timed, cores=cores, failed=failed, steps=k, wall=time.perf_counter() - t_wall)
