"""Synthetic randomised intersection scenarios (BASELINE.json configs 2, 4, 5).

A problem = one ego vehicle on one of the 12 A* routes of the reference's intersection
(main/envs/intersection.py, planned by main/lib/motion_primitive_search_modified.py), started at a random
path point with a lateral / heading / speed perturbation.  Problem k of a seed is the same whatever the
batch split, so a scenario sweep sharded over G ranks is the union of the 1-GPU sweep bit for bit
(`first` selects the shard)."""
import numpy as np

MAX_SPEED = 30.0 / 3.6      # main/lib/simulation.py:24


def make_workload(courses, n, seed=1235, first=0, tail_margin=200, lat_sigma=0.3, yaw_sigma=0.05):
    """Returns dict(route[n] i32, state[n,4] f64 (x,y,yaw,v), target_ind[n] i32, index0[n]).

    Each problem draws from its own counter-based stream (seed, global problem id) so shards compose."""
    route = np.empty(n, np.int32)
    state = np.empty((n, 4), np.float64)
    index0 = np.empty(n, np.int32)
    ids = np.arange(first, first + n, dtype=np.uint64)
    # one Philox stream per problem would be slow in Python; draw blocks of 8 doubles from a generator
    # keyed by (seed, block) instead: block = 4096 consecutive problem ids
    blk = 4096
    for b0 in range(int(first) // blk * blk, int(first) + n, blk):
        rng = np.random.Generator(np.random.Philox(key=[np.uint64(seed), np.uint64(b0 // blk)]))
        u = rng.random((blk, 3))
        g = rng.standard_normal((blk, 2))
        lo, hi = max(b0, first), min(b0 + blk, first + n)
        sl_src = slice(lo - b0, hi - b0)
        sl_dst = slice(lo - first, hi - first)
        r = np.minimum((u[sl_src, 0] * courses.R).astype(np.int32), courses.R - 1)
        nc = courses.length[r]
        i0 = np.minimum((u[sl_src, 1] * np.maximum(nc - tail_margin, 1)).astype(np.int32), nc - 1)
        px, py, pyaw = courses.packed[r, 0, i0], courses.packed[r, 1, i0], courses.packed[r, 2, i0]
        lat = lat_sigma * g[sl_src, 0]
        route[sl_dst] = r
        index0[sl_dst] = i0
        state[sl_dst, 0] = px - np.sin(pyaw) * lat
        state[sl_dst, 1] = py + np.cos(pyaw) * lat
        state[sl_dst, 2] = pyaw + yaw_sigma * g[sl_src, 1]
        state[sl_dst, 3] = u[sl_src, 2] * MAX_SPEED
    del ids
    return dict(route=route, state=state, target_ind=np.maximum(index0 - 3, 0).astype(np.int32), index0=index0)


def host_plant_step(state, delta, accel, L, dt, max_steer=np.deg2rad(45.0), max_speed=MAX_SPEED, min_speed=-5.0):
    """Simulation.step (main/lib/simulation.py:35-47, main/bicycle/main.py:28-41) vectorised over [B] on the
    host -- the caller-side plant of the e2e loop (in the reference the scenario script owns the simulation)."""
    x, y, yaw, v = state[:, 0], state[:, 1], state[:, 2], state[:, 3]
    d = np.clip(delta, -max_steer, max_steer)
    out = np.empty_like(state)
    out[:, 0] = x + (v * np.cos(yaw)) * dt
    out[:, 1] = y + (v * np.sin(yaw)) * dt
    out[:, 2] = yaw + ((v / L) * np.tan(d)) * dt
    out[:, 3] = np.clip(v + accel * dt, min_speed, max_speed)
    return out


def make_multi_agent_workload(courses, n_scenarios, vehicles=8, seed=1236, first=0, lane_span=15.0, min_gap=6.0,
                              v_max=5.0):
    """BASELINE.json configs[2]: per scenario `vehicles` egos on routes drawn from the 12
    (start lane, turn) pairs, at most 3 per start lane, placed on the approach lane at offsets in [0, lane_span]
    with gaps >= min_gap, initial speed U[0, v_max].  Ego k of scenario s is problem s*vehicles + k.
    Assumes route id = 3*lane + turn (the order of load_intersection_routes)."""
    n = n_scenarios * vehicles
    route = np.empty(n, np.int32)
    state = np.empty((n, 4))
    index0 = np.empty(n, np.int32)
    lanes = courses.R // 3
    assert vehicles <= 3 * lanes
    slots = np.repeat(np.arange(lanes), 3)
    for s in range(n_scenarios):
        rng = np.random.Generator(np.random.Philox(key=[np.uint64(seed), np.uint64(first + s)]))
        lane_of = rng.permutation(slots)[:vehicles]
        turn = rng.integers(0, 3, vehicles)
        off = np.zeros(vehicles)
        for ln in range(lanes):
            idx = np.flatnonzero(lane_of == ln)
            k = len(idx)
            if k:
                base = np.sort(rng.uniform(0.0, lane_span - min_gap * (k - 1), k))
                off[idx] = base + min_gap * np.arange(k)
        r = (3 * lane_of + turn).astype(np.int32)
        i0 = np.rint(off / courses.dl[r]).astype(np.int32)
        sl = slice(s * vehicles, (s + 1) * vehicles)
        route[sl], index0[sl] = r, i0
        state[sl, 0], state[sl, 1], state[sl, 2] = courses.packed[r, 0, i0], courses.packed[r, 1, i0], courses.packed[r, 2, i0]
        state[sl, 3] = rng.uniform(0.0, v_max, vehicles)
    return dict(route=route, state=state, target_ind=np.maximum(index0 - 3, 0).astype(np.int32), index0=index0)


def make_scripted_workload(courses, n, seed=1237, first=0, max_cars=3, v_max=4.0):
    """ego vehicles starting at the head of a random route (speed U[0, v_max]) with 0..max_cars
    scripted crossing cars each -- the constructor arguments of the reference's MovingObstacleTIntersection
    (main/lib/moving_obstacles.py:165-196; the reference scenario uses two, main/scenarios/mpc_intersection.py:41-44):
    direction +-1, turning or straight, speed U[3, 9] m/s, start offset 0 or U[0.2, 3] s.
    Returns dict(route, state, target_ind, scripts [n, 8, 4] with NaN rows past each ego's count)."""
    route = np.empty(n, np.int32)
    state = np.empty((n, 4))
    scripts = np.full((n, 8, 4), np.nan)
    for k in range(n):
        rng = np.random.Generator(np.random.Philox(key=[np.uint64(seed), np.uint64(first + k)]))
        r = int(rng.integers(0, courses.R))
        route[k] = r
        state[k] = (courses.packed[r, 0, 0], courses.packed[r, 1, 0], courses.packed[r, 2, 0], rng.uniform(0.0, v_max))
        for o in range(int(rng.integers(0, max_cars + 1))):
            scripts[k, o] = (float(rng.choice([-1, 1])), float(rng.integers(0, 2)), rng.uniform(3.0, 9.0),
                             float(rng.choice([0.0, rng.uniform(0.2, 3.0)])))
    return dict(route=route, state=state, target_ind=np.zeros(n, np.int32), scripts=scripts)
