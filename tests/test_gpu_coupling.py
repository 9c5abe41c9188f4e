"""Row A9 on the GPU (mpcb_coupling through the C ABI) against the vectors recorded from the reference's own
functions and against the numpy oracle on fresh scenes.  Integer outputs and flags must be bit-exact."""
import numpy as np
import pytest

import collision_oracle as co
from conftest import golden, load_routes

pytestmark = pytest.mark.gpu
CARS = [(2.86, 2.0, 3.5), (4.0, 2.04, 4.84)]


def _session(car=0, T=20):
    from mpc_for_av_at_intersection_b200 import CourseSet, MpcConfig, MpcSession
    L, w, ln = CARS[car]
    cfg = MpcConfig.from_json(T=T, wheelbase=L, car_width=w, car_length=ln)
    return MpcSession(cfg, CourseSet(load_routes(False)), 1024), cfg


def _pack_obs(obs_list):
    from mpc_for_av_at_intersection_b200 import _lib
    B = len(obs_list)
    obs = np.zeros((B, _lib.MAX_OBSTACLES, 6))
    n = np.zeros(B, np.int32)
    for b, o in enumerate(obs_list):
        n[b] = len(o)
        obs[b, :len(o)] = o
    return obs, n


@pytest.mark.parametrize("car", [0, 1])
def test_random_scenes_match_reference(car, routes_smooth):
    g = golden("collision_random.npz")
    sel = np.flatnonzero(g["car"] == car)
    ses, cfg = _session(car)
    route = g["route"][sel].astype(np.int32)
    state = np.zeros((len(sel), 4))
    for i, k in enumerate(sel):      # the ego sits exactly on its path point: the nearest-index update keeps agent_idx
        full = routes_smooth[int(g["route"][k])]
        j = int(g["agent_idx"][k])
        state[i] = (full[j, 0], full[j, 1], full[j, 2], g["v"][k])
    agent_idx = g["agent_idx"][sel].astype(np.int32)
    # previous cut = agent_idx + 1: the path had collapsed onto the ego's point, so the reference skips the
    # nearest-index update (mpc_intersection.py:103) and the recorded agent_idx is used as is
    ncourse = (agent_idx + 1).astype(np.int32)
    obs, n_obs = _pack_obs([g["obs"][k][:int(g["n_obs"][k])] for k in sel])
    out = ses.coupling(state, route, agent_idx, ncourse, obs, n_obs)
    assert (out["flags"] == 0).all()
    keep = np.ones(len(sel), bool)
    assert np.array_equal(agent_idx[keep], g["agent_idx"][sel][keep])
    assert np.array_equal(out["n_res"][keep], g["n_res"][sel][keep])
    assert np.array_equal(out["collided"][keep], g["collided"][sel][keep])
    assert np.array_equal(out["coll_idx"][keep], g["coll_idx"][sel][keep])
    assert np.array_equal(ncourse[keep], g["cutoff_idx"][sel][keep])
    hit = keep & (out["collided"] == 1)
    assert np.array_equal(out["coll_xy"][hit, 0], g["coll_x"][sel][hit]) and np.array_equal(out["coll_xy"][hit, 1], g["coll_y"][sel][hit])
    assert np.isnan(out["coll_xy"][keep & (out["collided"] == 0)]).all()
    print(f"car {car}: {keep.sum()} scenes, {int(out['collided'][keep].sum())} collisions, all integer outputs identical")


@pytest.mark.parametrize("T", [13, 20])
def test_scripted_obstacle_loop_replay(T, routes_smooth):
    """Every control step of the reference's obstacle scenario (mpc_intersection.py with two scripted cars)."""
    g = golden(f"closed_loop_T{T}.npz")
    sel = np.flatnonzero(g["loop"] == 12)
    ses, cfg = _session(0, T)
    B = len(sel)
    state = np.ascontiguousarray(g["state"][sel])
    route = g["route"][sel].astype(np.int32)
    prev_idx = np.concatenate([[0], g["agent_idx"][sel][:-1]]).astype(np.int32)      # traj_agent_idx before the update
    prev_nc = np.concatenate([[0], g["cutoff_idx"][sel][:-1]]).astype(np.int32)      # len(tmp_trajectory) of the previous step
    obs, n_obs = _pack_obs([g["obs"][k] for k in sel])
    agent_idx, ncourse = prev_idx.copy(), prev_nc.copy()
    out = ses.coupling(state, route, agent_idx, ncourse, obs, n_obs)
    assert (out["flags"] == 0).all()
    assert np.array_equal(agent_idx, g["agent_idx"][sel])
    assert np.array_equal(out["n_res"], g["n_res"][sel])
    assert np.array_equal(out["collided"], g["collided"][sel])
    assert np.array_equal(out["coll_idx"], g["coll_idx"][sel])
    assert np.array_equal(ncourse, g["cutoff_idx"][sel])
    assert np.array_equal(ncourse, g["ncourse"][sel])


def test_against_oracle_on_fresh_scenes(routes_smooth):
    rng = np.random.default_rng(11)
    ses, cfg = _session(0)
    B = 600
    route = rng.integers(0, 12, B).astype(np.int32)
    agent_idx = np.array([rng.integers(0, len(routes_smooth[r]) - 3) for r in route], np.int32)
    state = np.zeros((B, 4))
    obs_list = []
    for b in range(B):
        full = routes_smooth[route[b]]
        j = agent_idx[b]
        state[b] = (full[j, 0], full[j, 1], full[j, 2], rng.uniform(0, 8.3))
        n = int(rng.integers(0, 8))
        o = np.zeros((n, 6))
        for i in range(n):
            jj = min(len(full) - 1, j + int(rng.integers(0, 500)))
            o[i] = (full[jj, 0] + rng.normal(0, 4), full[jj, 1] + rng.normal(0, 4), rng.uniform(0, 9), rng.uniform(-3, 3),
                    rng.choice([0.0, rng.uniform(-1, 1)]), rng.choice([0.0, rng.uniform(-0.4, 0.4)]))
        obs_list.append(o)
    obs, n_obs = _pack_obs(obs_list)
    want_idx = agent_idx.copy()
    ncourse = (agent_idx + 1).astype(np.int32)
    out = ses.coupling(state, route, agent_idx, ncourse, obs, n_obs)
    assert np.array_equal(agent_idx, want_idx)
    offs, radius = co.circle_offsets(*CARS[0])
    for b in range(B):
        ref = co.coupling_step(routes_smooth[route[b]], int(want_idx[b]), float(state[b, 3]), obs_list[b], offs, radius, 2.86)
        got = (out["n_res"][b], out["collided"][b], out["coll_idx"][b], ncourse[b])
        assert got == (ref["n_res"], ref["collided"], ref["coll_idx"], ref["cutoff_idx"]), (b, got, ref)


def test_multi_agent_closed_loop_matches_oracle(routes_smooth):
    """BASELINE configs[2] spec: 8 interacting vehicles per scenario, every ego cut by the predicted
    motion of the others, Jacobi order.  Device loop (2 launches per control step) vs the CPU composition of the
    reference functions (oracle/collision_oracle.multi_agent_closed_loop): cut lengths and done flags bit-exact
    every step, states within 1e-6 (the first 30 steps are before any sensitive turn segment)."""
    from mpc_for_av_at_intersection_b200.scenarios import make_multi_agent_workload
    ses, cfg = _session(0, 20)
    V, S, K = 8, 2, 30
    wl = make_multi_agent_workload(ses.courses, S, V, seed=21)
    ses.begin(wl["route"], target_ind=wl["target_ind"])
    ses.store("state", wl["state"])
    offs, radius = co.circle_offsets(*CARS[0])
    refs = [co.multi_agent_closed_loop(routes_smooth, wl["route"][s * V:(s + 1) * V], wl["state"][s * V:(s + 1) * V],
                                       wl["target_ind"][s * V:(s + 1) * V], K, 20, offs, radius, 2.86) for s in range(S)]
    cuts = 0
    worst = 0.0
    state_ref = wl["state"].copy()
    for k in range(K):
        ses.multi_agent_dev(1, V)
        nc, st, done = ses.fetch("ncourse"), ses.fetch("state"), ses.fetch("done")
        for s in range(S):
            hist, done_ref, nclog = refs[s]
            sl = slice(s * V, (s + 1) * V)
            ran = ~np.isnan(hist[k, :, 0])
            state_ref[sl][ran] = hist[k, ran, :4]
            want_nc = np.where(nclog[k] > 0, nclog[k], ses.courses.length[wl["route"][sl]])
            assert np.array_equal(nc[sl], want_nc), (k, s, nc[sl], want_nc)
            cuts += int((nc[sl] < ses.courses.length[wl["route"][sl]]).sum())
            worst = max(worst, np.abs(st[sl] - state_ref[sl]).max())
    assert cuts > 20, "the scenarios must actually interact"
    print(f"multi-agent: {S}x{V} vehicles, {K} steps, {cuts} path cuts, max|dstate|={worst:.2e}")
    assert worst < 1e-6


REF_SCRIPTS = [(1, False, 25 / 3.6, 2.), (-1, True, 25 / 3.6, 4.)]        # main/scenarios/mpc_intersection.py:41-44


@pytest.mark.parametrize("T", [13, 20])
def test_scripted_loop_on_device_matches_reference(T, routes_smooth):
    """the reference's whole loop body WITH its two scripted cars (mpc_intersection.py:95-159) run on the
    device without host round trips (mpcb_session_scripted_loop_dev, two launches per control step), against the
    loop the unmodified reference recorded: steps to goal and every cut length bit-exact, the obstacle rows `o.get()`
    returned to 1e-12, positions within 1e-3 m over the rollout (1e-6 over the first 45 steps: left-turn route)."""
    import torch
    from mpc_for_av_at_intersection_b200 import _lib
    g = golden(f"closed_loop_T{T}.npz")
    sel = np.flatnonzero(g["loop"] == 12)
    n = len(sel)
    ses, cfg = _session(0, T)
    B = 3                                                   # replicas: every ego owns its copy of the cross traffic
    ses.begin(np.full(B, int(g["route"][sel[0]]), np.int32))
    ses.store("state", np.repeat(g["state"][sel[0]][None], B, 0))
    ses.set_scripted_obstacles([REF_SCRIPTS] * B)
    K = n + 3
    H = torch.full((K, B, 8), float("nan"), dtype=torch.float64, device="cuda")
    OH = torch.full((K, B, _lib.MAX_OBSTACLES, 6), float("nan"), dtype=torch.float64, device="cuda")
    before = ses.launches
    # step by step for the cut log, then compare the histories
    nclog = np.zeros((K, B), int)
    for k in range(K):
        ses.scripted_loop_dev(1, H[k].data_ptr(), OH[k].data_ptr())
        nclog[k] = ses.fetch("ncourse")
    torch.cuda.synchronize()
    assert ses.launches - before == 2 * K
    H, OH = H.cpu().numpy(), OH.cpu().numpy()
    assert np.array_equal(ses.fetch("done"), np.ones(B, np.int32))
    off = g["history_off"]
    ref = g["history"][off[12]:off[13]][1:n + 1]
    for b in range(B):
        valid = ~np.isnan(H[:, b, 0])
        assert valid.sum() == n and valid[:n].all(), (valid.sum(), n)
        assert np.array_equal(nclog[:n, b], g["ncourse"][sel])
        assert np.abs(OH[:n, b, :2] - g["obs"][sel]).max() < 1e-12
        ep = np.abs(H[:n, b, :2] - ref[:, :2]).max(axis=1)
        assert ep.max() < 1e-3 and ep[:45].max() < 1e-6, (ep.max(), ep[:45].max())
        assert np.abs(H[:45, b, 4:6] - ref[:45, 4:6]).max() < 1e-6
    print(f"T={T}: scripted loop on device, {n} steps, cuts bit-exact, max|dpos|={np.abs(H[:n, :, :2] - ref[None, :, :2].transpose(1, 0, 2)).max():.2e}")


def test_scripted_loop_random_scripts_vs_oracle_and_graph(routes_smooth):
    """Randomised cross traffic (direction, turning, speed, start offset; 0-3 cars per ego) on all 12 routes: the device
    loop captured into ONE CUDA graph (30 control steps = 60 kernel nodes, replayed by a single launch) against the
    composed numpy oracle -- cut lengths bit-exact every step, states within 1e-6."""
    import torch
    T, K = 20, 30
    rng = np.random.default_rng(5)
    ses, cfg = _session(0, T)
    B = 24
    route = (np.arange(B) % 12).astype(np.int32)
    scripts = []
    for b in range(B):
        scripts.append([(int(rng.choice([-1, 1])), bool(rng.integers(0, 2)), float(rng.uniform(3, 9)),
                         float(rng.choice([0.0, rng.uniform(0.2, 3.0)]))) for _ in range(int(rng.integers(0, 4)))])
    state0 = np.stack([np.array([routes_smooth[r][0, 0], routes_smooth[r][0, 1], routes_smooth[r][0, 2], rng.uniform(0, 4)])
                       for r in route])
    ses.begin(route)
    ses.store("state", state0)
    ses.set_scripted_obstacles(scripts)
    H = torch.full((K, B, 8), float("nan"), dtype=torch.float64, device="cuda")
    NC = torch.zeros((K, B), dtype=torch.int32, device="cuda")
    nc_view = torch.as_tensor(ses.device_view("ncourse"), device="cuda")
    stream = torch.cuda.Stream()
    graph = torch.cuda.CUDAGraph()
    # warm-up launch outside the capture (function attributes are set on first use), then reset the session
    ses.scripted_loop_dev(1)
    torch.cuda.synchronize()
    ses.begin(route)
    ses.store("state", state0)
    ses.set_scripted_obstacles(scripts)
    with torch.cuda.graph(graph, stream=stream):
        for k in range(K):
            ses.scripted_loop_dev(1, H[k].data_ptr(), 0, torch.cuda.current_stream().cuda_stream)
            NC[k].copy_(nc_view)
    before = ses.launches
    graph.replay()
    torch.cuda.synchronize()
    assert ses.launches == before, "the replay is one graph launch: no per-kernel host launches"
    H, NC = H.cpu().numpy(), NC.cpu().numpy()
    offs, radius = co.circle_offsets(*CARS[0])
    cuts, worst = 0, 0.0
    for b in range(B):
        full = routes_smooth[route[b]]
        hist, nclog, _, n = co.scripted_closed_loop(full, state0[b], scripts[b], K, T, offs, radius, 2.86)
        assert n == K
        assert np.array_equal(NC[:, b], nclog), (b, NC[:, b], nclog)
        cuts += int((nclog < len(full)).sum())
        worst = max(worst, np.abs(H[:, b, :6] - hist[:, :6]).max())
    assert cuts > 30, "the cross traffic must actually cut paths"
    print(f"scripted loop in one CUDA graph: {B} egos x {K} steps, {cuts} cut steps, max|dstate,u|={worst:.2e}")
    assert worst < 1e-6
