"""TEST INFRASTRUCTURE -- numpy restatement of the inter-vehicle coupling step.

Reference sequence per ego per control step (main/scenarios/mpc_intersection.py:100-139):
  nearest index on the full path -> ego "max-accel" prediction by resample_curve -> constant-input prediction of
  every obstacle (35 frames) -> all-pairs circle check over +-frame_window time offsets -> first hit -> position
  on the un-resampled path -> cutoff index.
Each function cites the reference lines it follows.  Pinned against tests/golden/collision_*.npz, which
oracle/gen_golden_collision.py records by calling the reference's OWN functions.  All outputs that matter are
integers/flags (bit-exact class); the float comparisons keep numpy's operation order (norm = sqrt(dx*dx + dy*dy)).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module."""
import math

import numpy as np

MAX_SPEED = 30.0 / 3.6


def circle_offsets(wheelbase, width, length, skip_back=False):
    """main/lib/car_dimensions.py:57-79: circle centres along the body axis, radius = width / sqrt(2)."""
    along = length / 2 - width / 2
    c = wheelbase / 2
    offs = [c + along] if skip_back else [c + along, c - along]
    return np.array(offs), width / (2 ** .5)


def predict_obstacle(x, y, v, yaw, a, steer, L, dt, horizon):
    """main/lib/moving_obstacles_prediction.py:21-47: x,y with the old v/yaw, then v, then yaw with the NEW v.
    Returns [n,3] (x, y, yaw), n = len(np.arange(0, horizon, dt))."""
    n = len(np.arange(0, horizon, dt))
    out = np.zeros((n, 3))
    for i in range(n):
        x += v * math.cos(yaw) * dt
        y += v * math.sin(yaw) * dt
        v += a * dt
        yaw += (v / L) * math.tan(steer) * dt
        out[i] = (x, y, yaw)
    return out


def resample_mask(points, dl):
    """main/lib/trajectories.py:58-86 -> boolean keep-mask (first and last always kept)."""
    n = len(points)
    d = points[1:, :2] - points[:-1, :2]
    step = np.concatenate([[0.0], np.sqrt(d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1])])
    q = np.floor(np.cumsum(step) / dl).astype(np.int64)
    keep = np.ones(n, bool)
    keep[1:] = (q[1:] - q[:-1]) >= 1
    keep[-1] = True
    return keep


def ego_prediction(path, v, dt, max_accel):
    """main/scenarios/mpc_intersection.py:110-116: spacing dl_k = dt*min(v + max_accel*(k+1), MAX_SPEED) per POINT."""
    if v < MAX_SPEED:
        dl = np.cumsum(np.zeros(len(path)) + max_accel) + v
        dl = dt * np.minimum(dl, MAX_SPEED)
    else:
        dl = dt * MAX_SPEED
    return path[resample_mask(path, dl)]


def _circle_points(traj, offs):
    """main/lib/trajectories.py:11-55 with y_offset = 0 -> [n, n_circles, 2]."""
    th = traj[:, 2]
    out = np.empty((len(traj), len(offs), 2))
    for c, xo in enumerate(offs):
        out[:, c, 0] = (np.cos(th) * xo - np.sin(th) * 0.0) + traj[:, 0]
        out[:, c, 1] = (np.sin(th) * xo + np.cos(th) * 0.0) + traj[:, 1]
    return out


def check_collision(offs, radius, traj_agent, path_detailed, trajs_obs, frame_window):
    """main/lib/collision_avoidance.py:66-104.  Returns None or dict(x, y, idx, row, frame).

    Row order of the reference's flattened mask: frame, agent circle, obstacle, time offset (-w..w), obstacle circle."""
    if len(trajs_obs) == 0:
        return None
    nc = len(offs)
    n_it = max(len(traj_agent), max(len(t) for t in trajs_obs))
    thr = 2 * radius
    fa = np.minimum(np.arange(n_it), len(traj_agent) - 1)                    # _pad_trajectory: repeat the last row
    ag = _circle_points(traj_agent, offs)[fa]                                  # [n_it, nc, 2]
    offsets = np.arange(-frame_window, frame_window + 1)
    obs = np.empty((n_it, len(trajs_obs), len(offsets), nc, 2))
    for o, tr in enumerate(trajs_obs):
        cc = _circle_points(tr, offs)
        g = np.minimum(np.arange(n_it), len(tr) - 1)
        for k, off in enumerate(offsets):                                      # _offset_trajectories_by_frames
            obs[:, o, k] = cc[np.clip(g - off, 0, len(tr) - 1)]
    dx = ag[:, :, None, None, None, 0] - obs[:, None, :, :, :, 0]
    dy = ag[:, :, None, None, None, 1] - obs[:, None, :, :, :, 1]
    mask = (np.sqrt(dx * dx + dy * dy) <= thr).reshape(-1)
    row = int(np.argmax(mask))
    if not mask[row]:
        return None
    per_frame = nc * len(trajs_obs) * len(offsets) * nc
    f, r = divmod(row, per_frame)
    r %= len(trajs_obs) * len(offsets) * nc
    o, r = divmod(r, len(offsets) * nc)
    k, c = divmod(r, nc)
    opos = obs[f, o, k, c]
    pc = _circle_points(path_detailed, offs)                                   # [Np, nc, 2]
    flat = np.concatenate([pc[:, c2] for c2 in range(nc)])                     # circle-major, as the reference
    ddx, ddy = opos[0] - flat[:, 0], opos[1] - flat[:, 1]
    m2 = np.sqrt(ddx * ddx + ddy * ddy) <= thr
    idx = int(np.argmax(m2)) % len(path_detailed)                              # no hit -> 0 (reference behaviour)
    return dict(x=float(path_detailed[idx, 0]), y=float(path_detailed[idx, 1]), idx=idx, row=row, frame=f)


def cutoff_index(full, x, y, radius=0.001):
    """main/lib/collision_avoidance.py:107-119: first path point within `radius`; -1 where the reference returns the
    array itself (latent bug, unreachable when (x, y) is a path point)."""
    dx, dy = full[:, 0] - x, full[:, 1] - y
    m = np.sqrt(dx * dx + dy * dy) <= radius
    j = int(np.argmax(m))
    return j if m[j] else -1


def coupling_step(full, agent_idx, v, obstacles, offs, radius, L, dt=0.2, max_accel=2.0, horizon=7.0,
                  frame_window=20, extra_margin=None):
    """The loop body main/scenarios/mpc_intersection.py:108-136 after the nearest-index update.
    obstacles: [n_obs, 6] rows (x, y, v, yaw, a, steer) as MovingObstacle*.get() returns them.
    Returns dict(n_res, collided, coll_x, coll_y, coll_idx, row, cutoff_idx)."""
    dl = float(np.linalg.norm(full[0, :2] - full[1, :2]))
    if extra_margin is None:
        extra_margin = 4 * int(math.ceil(radius / dl))
    path = full[agent_idx:]
    res = ego_prediction(path, v, dt, max_accel)
    trajs = [predict_obstacle(*o, L, dt, horizon) for o in obstacles]
    hit = check_collision(offs, radius, res, path, trajs, frame_window)
    out = dict(n_res=len(res), collided=0, coll_x=np.nan, coll_y=np.nan, coll_idx=-1, row=-1, cutoff_idx=len(full))
    if hit is not None:
        j = cutoff_index(full, hit["x"], hit["y"])
        out.update(collided=1, coll_x=hit["x"], coll_y=hit["y"], coll_idx=hit["idx"], row=hit["row"],
                   cutoff_idx=max(agent_idx + 1, j - extra_margin))
    return out


def agent_index_update(full, agent_idx, ncourse_prev, x, y):
    """main/scenarios/mpc_intersection.py:103-105: skip the update when the cut path has collapsed onto the
    ego's point (ncourse_prev = len(tmp_trajectory) of the previous step, 0 = None).  -1 = the reference raises."""
    import mpc_oracle as mo
    if ncourse_prev > 0 and not np.any(full[agent_idx, :3] != full[ncourse_prev - 1, :3]):
        return agent_idx
    return mo.nearest_index_in_direction(x, y, full[:, 0], full[:, 1], agent_idx)


def multi_agent_closed_loop(routes_smooth, route, state0, target0, nsteps, T, offs, radius, L, dt=0.2):
    """SURVEY.md section 8a, A9 spec for config 3: every ego takes the OTHER egos of its scenario as obstacles
    (constant speed, zero steer), Jacobi order.  `route/state0/target0` describe ONE scenario of V vehicles.
    Returns history [nsteps, V, 8] (NaN rows once an ego is done), done[V], ncourse log [nsteps, V]."""
    import mpc_oracle as mo
    V = len(route)
    cfg = mo.Cfg.from_dict({"T": T})
    mpcs = []
    for j in range(V):
        full = routes_smooth[route[j]]
        m = mo.OracleMPC(full[:, 0].copy(), full[:, 1].copy(), full[:, 2].copy(),
                         float(np.linalg.norm(full[0, :2] - full[1, :2])), L, dt, cfg)
        m.target_ind = int(target0[j])
        mpcs.append(m)
    state = np.array(state0, float)
    agent_idx = np.zeros(V, int)
    ncourse = np.zeros(V, int)
    done = np.zeros(V, int)
    hist = np.full((nsteps, V, 8), np.nan)
    nclog = np.zeros((nsteps, V), int)
    for k in range(nsteps):
        new_state = state.copy()
        for j in range(V):
            if done[j]:
                nclog[k, j] = ncourse[j]
                continue
            full = routes_smooth[route[j]]
            m = mpcs[j]
            if m.is_goal(state[j]):                 # loop head, with the path as cut by the PREVIOUS step (:97)
                done[j] = 1
                nclog[k, j] = ncourse[j]
                continue
            ai = agent_index_update(full, int(agent_idx[j]), int(ncourse[j]), state[j, 0], state[j, 1])
            assert ai >= 0
            agent_idx[j] = ai
            others = [(state[i, 0], state[i, 1], state[i, 3], state[i, 2], 0.0, 0.0) for i in range(V) if i != j]
            c = coupling_step(full, ai, state[j, 3], others, offs, radius, L, dt)
            ncourse[j] = c["cutoff_idx"]
            nclog[k, j] = ncourse[j]
            m.set_trajectory_fromarray(full[:ncourse[j]])
            di, ai_ = m.step(state[j])
            dev = m.get_current_xref_deviation() if (m.ox is not None and m.target_ind < len(m.cx)) else np.nan
            new_state[j] = mo.plant_step(state[j, 0], state[j, 1], state[j, 2], state[j, 3], ai_, di, L, dt)
            hist[k, j] = (*new_state[j], di, ai_, dev, m.status)
        state = new_state
    return hist, done, nclog


class ScriptedObstacle:
    """main/lib/moving_obstacles.py:165-231 (`MovingObstacleTIntersection`) with the Bicycle model of
    main/bicycle/main.py:28-41: a scripted car crossing the junction, independent of the ego."""

    def __init__(self, direction, turning, speed, offset, L, dt):
        self.direction = 1 if direction >= 0 else -1                      # :177
        self.turning, self.speed, self.L, self.dt = bool(turning), speed, L, dt
        self.offset = None if offset is None else offset if offset > 0 else None     # :181
        self.counter = 0
        if self.direction == 1:                                           # :186-196
            self.xc, self.yc, self.theta, self.x_turn = -30, -3, 0, -10
        else:
            self.xc, self.yc, self.theta, self.x_turn = 30, 3, np.pi, 12

    @property
    def steering_angle(self):                                            # :201-216
        if self.turning is not True:
            return 0.
        if self.direction == 1:
            if self.xc >= self.x_turn and self.theta > (-np.pi / 2):
                return -0.38
        else:
            if self.xc <= self.x_turn and self.theta < (3 * np.pi / 2):
                return 0.19
        return 0.

    @property
    def forward_velocity(self):                                          # :218-224
        if self.offset is None or self.counter > (self.offset / self.dt):
            return self.speed
        return 0

    def step(self):                                                      # :225-228, bicycle/main.py:34-41
        delta, v = self.steering_angle, self.forward_velocity
        xd = v * np.cos(self.theta)
        yd = v * np.sin(self.theta)
        td = (v / self.L) * np.tan(delta)
        self.xc += xd * self.dt
        self.yc += yd * self.dt
        self.theta += td * self.dt
        self.counter += 1

    def get(self):                                                       # :230-232
        return self.xc, self.yc, self.forward_velocity, self.theta, 0.0, self.steering_angle


def scripted_closed_loop(full, state0, scripts, nsteps, T, offs, radius, L, dt=0.2, target0=0):
    """The loop body of main/scenarios/mpc_intersection.py:95-159 for one ego with scripted cross traffic
    (`scripts`: (direction, turning, speed, offset) tuples).  Returns history [nsteps, 8] (x, y, yaw, v after the plant
    step, delta, a, xref_deviation, status; NaN rows once the goal is reached), the cut log [nsteps] (0 once done), the
    obstacle rows [nsteps, n_obs, 6] that `get()` returned, and the number of steps run."""
    import mpc_oracle as mo
    cfg = mo.Cfg.from_dict({"T": T})
    m = mo.OracleMPC(full[:, 0].copy(), full[:, 1].copy(), full[:, 2].copy(),
                     float(np.linalg.norm(full[0, :2] - full[1, :2])), L, dt, cfg)
    m.target_ind = int(target0)
    obstacles = [ScriptedObstacle(*s, L, dt) for s in scripts]
    state = np.array(state0, float)
    agent_idx, ncourse = 0, 0
    hist = np.full((nsteps, 8), np.nan)
    nclog = np.zeros(nsteps, int)
    obslog = np.full((nsteps, len(obstacles), 6), np.nan)
    for k in range(nsteps):
        if m.is_goal(state):                                             # :97
            return hist, nclog, obslog, k
        agent_idx = agent_index_update(full, agent_idx, ncourse, state[0], state[1])      # :103-105
        assert agent_idx >= 0
        rows = [o.get() for o in obstacles]                              # :118-121
        obslog[k] = np.array(rows, float).reshape(len(obstacles), 6)
        c = coupling_step(full, agent_idx, state[3], rows, offs, radius, L, dt)
        ncourse = c["cutoff_idx"]
        nclog[k] = ncourse
        m.set_trajectory_fromarray(full[:ncourse])                       # :139
        di, ai = m.step(state)                                           # :142
        for o in obstacles:                                              # :154-156
            o.step()
        dev = m.get_current_xref_deviation() if (m.ox is not None and m.target_ind < len(m.cx)) else np.nan
        state = np.array(mo.plant_step(state[0], state[1], state[2], state[3], ai, di, L, dt))   # :159
        hist[k] = (*state, di, ai, dev, m.status)
    return hist, nclog, obslog, nsteps
