"""Parity ON THE WORKLOADS THAT ARE BENCHMARKED: every control step of the closed loops that
bench.py / scripts/bench_multi_agent.py / scripts/sweep_config5.py time is replayed, on IDENTICAL per-step inputs, through
the independent plain C++ CPU restatement of main/lib/mpc.py (oracle/mpc_port.cpp, pinned against the goldens recorded
from the unmodified reference in tests/test_cpp_port.py).

What "failed" means in the reference: `_linear_mpc_control` returns None unless the solver status is OPTIMAL or
OPTIMAL_INACCURATE (main/lib/mpc.py:196-208) and `MPC.step` then applies `ai = MAX_DECEL`, keeps `di` and drops the warm
start (:294-297).  That status class (solved / failed) is a flag -- the bit-exact class: the kernel must never fail a
problem the CPU solver solves.

Bars: status class equal on every problem; target_ind, xref and cut lengths array_equal; plant states 1e-9; applied
controls |d(di, ai)| <= 1e-6 on 99.9 % of the solves and <= 1e-4 (the north_star bar) on all of them.  Why not 1e-6 on
all: where a bound is only WEAKLY active (multiplier ~1e-7) an interior point stops O(sqrt(mu)) inside it -- at mu = 1e-12
the kernel and the port both sit 2e-6 .. 2e-5 away from the vertex that the active-set-polished numpy oracle certifies
(profiles/r02_NOTES.md lists the worst cases replayed through it); ECOS at its 1e-8 tolerance is no closer."""
import os
import sys

import numpy as np
import pytest

from conftest import REPO, load_routes

pytestmark = pytest.mark.gpu

PRE = ("state", "oa", "odelta", "target_ind", "warm", "di_ai", "ncourse", "route")


def _setup(T, B):
    import mpc_port
    from mpc_for_av_at_intersection_b200 import CourseSet, MpcConfig, MpcSession, load_intersection_routes
    cfg = MpcConfig.from_json(T=T)
    cs = CourseSet(load_intersection_routes())
    return cfg, cs, MpcSession(cfg, cs, B), mpc_port.pack_config(cfg), mpc_port


def _port_replay(mpc_port, cfgvec, cs, pre, L, sel=None, want_traj=False):
    """MPC.step of the CPU port on the pre-step object state `pre` (copies); returns dict(status, target, warm, oa, od,
    di_ai [, xref])."""
    sel = np.arange(len(pre["state"])) if sel is None else np.asarray(sel)
    c = lambda a, dt: np.ascontiguousarray(a[sel]).astype(dt).copy()
    state, route, nc = c(pre["state"], np.float64), c(pre["route"], np.int32), c(pre["ncourse"], np.int32)
    tgt, warm = c(pre["target_ind"], np.int32), c(pre["warm"], np.int32)
    oa, od, di_ai = c(pre["oa"], np.float64), c(pre["odelta"], np.float64), c(pre["di_ai"], np.float64)
    out = mpc_port.step_batch(cfgvec, cs.packed, cs.dl, state, route, nc, np.full(len(sel), float(L)), tgt, warm, oa, od, di_ai,
                              want_traj=want_traj)
    return dict(status=out["status"], iters=out["iters"], target=tgt, warm=warm, oa=oa, od=od, di_ai=di_ai, xref=out["xref"])


def _dump(name, pre, idx, extra):
    """failing problems go to gpurun_out/ so that they can be replayed on a CPU-only box"""
    os.makedirs(os.path.join(REPO, "gpurun_out"), exist_ok=True)
    np.savez(os.path.join(REPO, "gpurun_out", name), **{k: pre[k][idx] for k in PRE}, **extra)


def _compare_step(tag, k, pre, post, ref, ran, cfg, stats):
    """post: the device's object state after the step; ref: the port's on the same inputs; ran: problems that solved"""
    g_solved, p_solved = post["status"][ran] <= 1, ref["status"] <= 1
    g_near, p_near = post["status"][ran] == 3, ref["status"] == 3
    assert np.array_equal(g_near, p_near), f"{tag} step {k}: nearest-index exception class differs"
    diff = np.flatnonzero(g_solved != p_solved)
    if len(diff):
        idx = np.flatnonzero(ran)[diff]
        _dump(f"status_mismatch_{tag}_step{k}.npz", pre, idx, dict(gpu_status=post["status"][idx], port_status=ref["status"][diff],
                                                                  gpu_iters=post["iters"][idx], port_iters=ref["iters"][diff]))
    kernel_only_fails = int((~g_solved & p_solved).sum())
    stats["kernel_fails_port_solves"] += kernel_only_fails
    stats["port_fails_kernel_solves"] += int((g_solved & ~p_solved).sum())
    stats["both_fail"] += int((~g_solved & ~p_solved & ~g_near).sum())
    stats["solves"] += int(ran.sum())
    both = g_solved & p_solved
    ok = ~g_near
    assert np.array_equal(post["target_ind"][ran][ok], ref["target"][ok]), f"{tag} step {k}: target_ind"
    assert np.array_equal(post["warm"][ran][both], ref["warm"][both])
    err = np.abs(post["di_ai"][ran][both] - ref["di_ai"][both]).max(axis=1) if both.any() else np.zeros(1)
    e = err.max()
    if e > max(stats["max_du0"], 2e-7):              # keep the worst case for a replay through the KKT-certified oracle
        w = np.flatnonzero(ran)[np.flatnonzero(both)[int(err.argmax())]]
        _dump(f"worst_du0_{tag}.npz", pre, [w], dict(gpu_di_ai=post["di_ai"][[w]], port_di_ai=ref["di_ai"][both][[int(err.argmax())]],
                                                    gpu_iters=post["iters"][[w]], step=np.array([k])))
    stats["max_du0"] = max(stats["max_du0"], float(e))
    stats["errs"].append(err)
    f = ~g_solved & ~p_solved & ok                     # the reference's fallback on both sides (mpc.py:294-297)
    assert np.all(post["di_ai"][ran][f, 1] == cfg.MAX_DECEL) and np.array_equal(post["di_ai"][ran][f, 0], pre["di_ai"][ran][f, 0])
    return kernel_only_fails


def _new_stats():
    return dict(kernel_fails_port_solves=0, port_fails_kernel_solves=0, both_fail=0, solves=0, max_du0=0.0, errs=[])


def _check_controls(stats):
    errs = np.concatenate(stats.pop("errs"))
    stats["p999_du0"] = float(np.percentile(errs, 99.9))
    return stats


def test_config2_as_benched_every_step_vs_port():
    """BASELINE.json configs[1] exactly as bench.py runs it: make_workload(seed=1235), B = 4096, T = 20, the first 25
    closed-loop steps from the same cold start (bench: 5 warm-up + 20 timed), fused solve + plant, goal test off."""
    from mpc_for_av_at_intersection_b200.scenarios import host_plant_step, make_workload
    T, B, K = 20, 4096, 25
    cfg, cs, ses, cfgvec, mpc_port = _setup(T, B)
    wl = make_workload(cs, B, seed=1235)
    ses.begin(wl["route"], target_ind=wl["target_ind"])
    ses.store("state", wl["state"])
    stats = _new_stats()
    ran = np.ones(B, bool)
    for k in range(K):
        pre = {n: ses.fetch(n) for n in PRE}
        ses.closed_loop_dev(1, 0, 0, goal_test=False)
        post = {n: ses.fetch(n) for n in ("status", "iters", "target_ind", "warm", "di_ai", "state", "xref")}
        ref = _port_replay(mpc_port, cfgvec, cs, pre, cfg.wheelbase, want_traj=True)
        _compare_step("cfg2", k, pre, post, ref, ran, cfg, stats)
        assert np.array_equal(post["xref"], ref["xref"]), "xref is gathered by integer indices: bit-exact"
        want = host_plant_step(pre["state"], post["di_ai"][:, 0], post["di_ai"][:, 1], cfg.wheelbase, cfg.dt)
        assert np.abs(post["state"] - want).max() < 1e-9
    print("config 2 as benched:", _check_controls(stats))
    assert stats["solves"] == B * K
    assert stats["kernel_fails_port_solves"] == 0 and stats["port_fails_kernel_solves"] == 0
    assert stats["p999_du0"] < 1e-6 and stats["max_du0"] < 1e-4


def test_config3_as_benched_every_step_vs_port(routes_smooth):
    """BASELINE.json configs[2]: 1024 scenarios x 8 interacting vehicles (make_multi_agent_workload, as
    scripts/bench_multi_agent.py), 40 control steps of the fused device loop (coupling launch + step launch), goal test
    on.  Every solve is replayed through the port on the cut path the coupling kernel chose; the cut itself is checked
    against the CPU composition of the reference's functions (oracle/collision_oracle.py) on a sample of scenarios."""
    import collision_oracle as co
    from mpc_for_av_at_intersection_b200.scenarios import make_multi_agent_workload
    T, S, V, K = 20, 1024, 8, 40
    B = S * V
    cfg, cs, ses, cfgvec, mpc_port = _setup(T, B)
    wl = make_multi_agent_workload(cs, S, V)
    ses.begin(wl["route"], target_ind=wl["target_ind"])
    ses.store("state", wl["state"])
    offs, radius = co.circle_offsets(2.86, 2.0, 3.5)
    stats = _new_stats()
    cuts = checked = 0
    rng = np.random.default_rng(3)
    for k in range(K):
        pre = {n: ses.fetch(n) for n in PRE + ("done", "agent_idx")}
        ses.multi_agent_dev(1, V)
        post = {n: ses.fetch(n) for n in ("status", "iters", "target_ind", "warm", "di_ai", "state", "ncourse", "done", "agent_idx")}
        ran = (pre["done"] == 0) & (post["done"] != 1)      # not finished before, not finished by this step's goal test
        assert ran.sum() > 0
        used = dict(pre, ncourse=post["ncourse"])          # the step solves on the path as cut by THIS step's coupling
        ref = _port_replay(mpc_port, cfgvec, cs, used, cfg.wheelbase, sel=np.flatnonzero(ran))
        _compare_step("cfg3", k, used, post, ref, ran, cfg, stats)
        cuts += int((post["ncourse"][ran] < cs.length[wl["route"]][ran]).sum())
        for s in rng.choice(S, 6, replace=False):          # cut lengths: bit-exact class
            for j in range(V):
                b = s * V + j
                if not ran[b]:
                    continue
                full = routes_smooth[wl["route"][b]]
                ai = co.agent_index_update(full, int(pre["agent_idx"][b]), int(pre["ncourse"][b]) if k else 0,
                                           pre["state"][b, 0], pre["state"][b, 1])
                others = [(pre["state"][i, 0], pre["state"][i, 1], pre["state"][i, 3], pre["state"][i, 2], 0.0, 0.0)
                          for i in range(s * V, (s + 1) * V) if i != b]
                c = co.coupling_step(full, ai, pre["state"][b, 3], others, offs, radius, cfg.wheelbase, cfg.dt)
                assert (post["agent_idx"][b], post["ncourse"][b]) == (ai, c["cutoff_idx"]), (k, b)
                checked += 1
    print("config 3 as benched:", _check_controls(stats), "cut steps", cuts, "coupling checks", checked)
    assert cuts > 1000, "the scenarios must actually interact"
    assert stats["kernel_fails_port_solves"] == 0 and stats["port_fails_kernel_solves"] == 0
    assert stats["p999_du0"] < 1e-6 and stats["max_du0"] < 1e-4


@pytest.mark.parametrize("T,B,K,L", [(10, 8192, 13, 2.86), (13, 8192, 13, 2.86), (40, 8192, 13, 2.86), (80, 4096, 25, 2.86),
                                     (80, 65536, 11, 4.0)])      # the last one: the sweep's largest cell, Prius wheelbase
def test_config5_long_and_short_horizons_never_fail_where_the_port_solves(T, B, K, L):
    """BASELINE.json configs[4] generator (cfg-2 problems at other horizons), >= 100 K solves per horizon from a cold
    start: every problem the kernel reports as failed is replayed through the port, which must fail it too (zero
    'kernel fails, port solves'); a sample of the solved ones is compared at 1e-6."""
    from mpc_for_av_at_intersection_b200.scenarios import make_workload
    cfg, cs, ses, cfgvec, mpc_port = _setup(T, B)
    wl = make_workload(cs, B, seed=1234 + 4)
    ses.begin(wl["route"], target_ind=wl["target_ind"], L=np.full(B, L))
    ses.store("state", wl["state"])
    rng = np.random.default_rng(T)
    failed = kernel_only = sampled = 0
    worst = 0.0
    errs = []
    for k in range(K):
        pre = {n: ses.fetch(n) for n in PRE}
        ses.closed_loop_dev(1, 0, 0, goal_test=False)
        post = {n: ses.fetch(n) for n in ("status", "iters", "di_ai", "target_ind")}
        bad = np.flatnonzero(post["status"] == 2)
        failed += len(bad)
        if len(bad):
            ref = _port_replay(mpc_port, cfgvec, cs, pre, L, sel=bad)
            wrong = bad[ref["status"] <= 1]
            if len(wrong):
                _dump(f"kernel_only_failures_T{T}_step{k}.npz", pre, wrong,
                      dict(gpu_iters=post["iters"][wrong], port_di_ai=ref["di_ai"][ref["status"] <= 1]))
            kernel_only += len(wrong)
        good = np.flatnonzero(post["status"] <= 1)
        sel = rng.choice(good, min(len(good), 48 if T >= 40 else 256), replace=False)
        ref = _port_replay(mpc_port, cfgvec, cs, pre, L, sel=sel)
        both = ref["status"] <= 1
        assert both.mean() > 0.98
        assert np.array_equal(post["target_ind"][sel], ref["target"])
        err = np.abs(post["di_ai"][sel][both] - ref["di_ai"][both]).max(axis=1)
        if err.max() > max(worst, 2e-7):
            w = sel[both][int(err.argmax())]
            _dump(f"worst_du0_T{T}.npz", pre, [w], dict(gpu_di_ai=post["di_ai"][[w]], port_di_ai=ref["di_ai"][both][[int(err.argmax())]],
                                                      gpu_iters=post["iters"][[w]], step=np.array([k])))
        worst = max(worst, float(err.max()))
        errs.append(err)
        sampled += len(sel)
    print(f"T={T}: {B * K} solves, {failed} failed on the device, {kernel_only} of them solved by the port; "
          f"{sampled} sampled solves max|du0|={worst:.2e} p99={np.percentile(np.concatenate(errs), 99):.2e}")
    assert kernel_only == 0
    assert np.percentile(np.concatenate(errs), 99) < 1e-6 and worst < 1e-4


def test_hard_cases_recorded_from_round_1_failures():
    """tests/golden/hard_cases.npz: the twelve bench-workload steps on which the round-1 kernels reported FAILED while the
    CPU port solved them (one config-2 step at v0 = MAX_SPEED, one T = 40 and ten T = 80 steps of vehicles far off their
    path) -- the flipped status flag of VERDICT r01 weak #1, plus the scripts/find_fail.py case -- and a thirteenth found by
    the round-2 soak sweeps (T = 80, 1 solve in 1.1 M: the port reports its inaccurate class after 42 iterations, the numpy
    oracle gives up; scripts/triage_sweep_failures.py).  The kernel must solve every one and land on the port's controls."""
    from mpc_for_av_at_intersection_b200 import CourseSet, MpcConfig, load_intersection_routes
    from mpc_for_av_at_intersection_b200.batched import BatchedMPC
    from conftest import golden
    g = golden("hard_cases.npz")
    cs = CourseSet(load_intersection_routes())
    for T in (20, 40, 80):
        sel = np.flatnonzero(g["T"] == T)
        eng = BatchedMPC(MpcConfig.from_json(T=T), cs, len(sel))
        c = lambda k, dt: np.ascontiguousarray(g[k][sel]).astype(dt)
        tgt, warm, di_ai = c("target_ind", np.int32), c("warm", np.int32), c("di_ai", np.float64)
        oa, od = np.ascontiguousarray(g["oa"][sel][:, :T]), np.ascontiguousarray(g["odelta"][sel][:, :T])
        out = eng.step_host(c("state", np.float64), c("route", np.int32), c("ncourse", np.int32), tgt, warm, oa, od, di_ai, want_traj=False)
        assert (g["port_status"][sel] <= 1).all()
        assert (out["status"] <= 1).all(), (T, out["status"], out["iters"])
        assert np.array_equal(tgt, g["port_target"][sel])
        err = np.abs(di_ai - g["port_di_ai"][sel]).max()
        print(f"hard cases T={T}: {len(sel)} problems, iterations {out['iters']} (port {g['port_iters'][sel]}), max|du0|={err:.2e}")
        assert err < 1e-4
        eng.close()
