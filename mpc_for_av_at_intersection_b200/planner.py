"""Batched motion-primitive A* planner: the step BEFORE the MPC loop in the reference --
`MotionPrimitiveSearch(scenario, car_dimensions, mps, margin=car_dimensions.radius).run()`
(main/scenarios/mpc_intersection.py:56-60, main/lib/motion_primitive_search_modified.py) -- with the same class names
and call shapes, running on the GPU through `mpcb_plan_routes_host` (one warp per search), plus `plan_batch` for many
scenarios per call.  There is no CPU fallback.

Host-side pieces (they run once per scenario / primitive library, like the reference's constructor work):
  * `Scenario`, `BoxObstacle`, `CircleObstacle` with `to_convex(margin)` -> half-plane rows (main/lib/scenario.py:7-13,
    main/lib/obstacles.py:46-151);
  * `intersection(turn_indicator, start_pos)`: the four-arm junction of main/envs/intersection.py:10-213 (same numbers,
    built arm by arm from one table);
  * `load_motion_primitives(version)`: the reference's primitive library (main/data/motion_primitives_*/*.pkl) shipped as
    arrays in data/motion_primitives.npz (written by oracle/gen_golden_astar.py from the unmodified pickles);
  * collision points of every primitive: `resample_curve(points, dl=radius)` then the car's circle centres
    (main/lib/motion_primitive_search_modified.py:35-52, main/lib/trajectories.py:11-86)."""
import ctypes as C
import math
import os
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from . import _lib

MAX_OBS, MAX_HP = 32, 8
PLAN_OK, PLAN_NO_SOLUTION, PLAN_OVERFLOW = 0, 1, 2
_DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "motion_primitives.npz")


# ---- obstacles and scenarios (main/lib/obstacles.py, main/lib/scenario.py) ------------------------------------------
class BoxObstacle:
    def __init__(self, xy_width, height, xy_center, hidden=False):
        self.xy_width, self.height, self.xy_center, self.hidden = tuple(xy_width), height, tuple(xy_center), hidden
        (cx, cy), (w, h) = xy_center, xy_width
        self.xy1 = cx - w / 2, cy - h / 2
        self.xy2 = cx + w / 2, cy + h / 2

    def to_convex(self, margin=0.):
        """obstacles.py:96-105: right, left, top, bottom."""
        (x1, y1), (x2, y2) = self.xy1, self.xy2
        return np.array([[1, 0, -(x2 + margin)], [-1, 0, x1 - margin], [0, 1, -(y2 + margin)], [0, -1, y1 - margin]])

    def distance_to_point(self, point):
        (x1, y1), (x2, y2), (x, y) = self.xy1, self.xy2, point
        dx, dy = max(x1 - x, 0, x - x2), max(y1 - y, 0, y - y2)
        return math.sqrt(dx * dx + dy * dy)


class CircleObstacle:
    def __init__(self, radius, height, xy_center, hidden=False):
        self.radius, self.height, self.xy_center, self.hidden = radius, height, tuple(xy_center), hidden

    def to_convex(self, margin=0.):
        """obstacles.py:135-151: the circumscribed octagon."""
        (cx, cy), r = self.xy_center, self.radius
        d, m2 = r * np.sqrt(2), 2 * margin                 # subtracted one after the other, as the reference does
        return np.array([[1, 0, -(cx + r + margin)], [-1, 0, cx - r - margin], [0, 1, -(cy + r + margin)],
                         [0, -1, cy - r - margin], [-1, 1, cx - cy - d - m2], [1, -1, -cx + cy - d - m2],
                         [-1, -1, cx + cy - d - m2], [1, 1, -cx - cy - d - m2]])

    def distance_to_point(self, point):
        return max(0, math.sqrt((self.xy_center[0] - point[0]) ** 2 + (self.xy_center[1] - point[1]) ** 2) - self.radius)


@dataclass
class Scenario:
    start: Tuple[float, float, float]
    goal_point: Tuple[float, float, float]
    goal_area: BoxObstacle
    allowed_goal_theta_difference: float
    obstacles: list


def intersection(turn_indicator: int, start_pos: int) -> Scenario:
    """The urban junction of main/envs/intersection.py:10-213.  start_pos 1 south, 2 west, 3 north, 4 east;
    turn_indicator 1 left, 2 straight, 3 right."""
    road, island, pavement, length, height, corner = 4, 2, 5, 30, 0.5, 6
    dc = corner + road + island                    # distance_center
    lane = island / 2 + road / 2                   # lane centre offset
    far = length / 2 + dc                          # centre of the arm boxes
    side = island / 2 + road + pavement / 2        # centre of the pavements
    half_pi = 0.5 * np.pi
    starts = {1: (lane, -30, half_pi), 2: (-30, -lane, 0), 3: (-lane, 30, -half_pi), 4: (30, lane, np.pi)}
    mid, gd = (island + road) / 2, 30
    west, north, east, south = (-gd, mid, -np.pi), (mid, gd, half_pi), (gd, -mid, 0), (-mid, -gd, -half_pi)
    goals = {1: (west, north, east), 2: (north, east, south), 3: (east, south, west), 4: (south, west, north)}
    goal = goals[start_pos][turn_indicator - 1]
    wide_x = (start_pos in (1, 3) and turn_indicator in (1, 3)) or (start_pos in (2, 4) and turn_indicator in (2, 4))
    goal_area = BoxObstacle((road * 1.8, road) if wide_x else (road, road * 1.8), height, (goal[0], goal[1]))
    arms = [(0, -1), (0, 1), (-1, 0), (1, 0)]      # south, north, west, east (the reference's listing order)

    def arm_box(ax, ay, across, w_across):
        """Box centred `across` off the arm's axis, `w_across` wide, `length` long, at the far end of arm (ax, ay)."""
        if ax == 0:
            return BoxObstacle((w_across, length), height, (across, ay * far))
        return BoxObstacle((length, w_across), height, (ax * far, across))

    obstacles = []
    for ax, ay in arms:                            # medians: box + round nose
        obstacles.append(arm_box(ax, ay, 0, island))
        obstacles.append(CircleObstacle(island / 2, height, (ax * dc, ay * dc)))
    for cx, cy in ((-1, -1), (-1, 1), (1, 1), (1, -1)):      # corners
        obstacles.append(CircleObstacle(dc - island / 2 - road, height, (cx * dc, cy * dc)))
    for ax, ay in ((0, -1), (-1, 0), (0, 1), (1, 0)):        # pavements: south, west, north, east
        obstacles.append(arm_box(ax, ay, -side, pavement))
        obstacles.append(arm_box(ax, ay, side, pavement))
    # hidden boxes that close the oncoming lanes (traffic rules), depending on where the car comes from (:149-203):
    # sign of the closed lane on the west, east, south, north arm
    closed = {1: (-1, 1, -1, -1), 2: (1, 1, 1, -1), 3: (-1, 1, 1, 1), 4: (-1, -1, 1, -1)}[start_pos]
    for (ax, ay), sgn in zip(((-1, 0), (1, 0), (0, -1), (0, 1)), closed):
        b = arm_box(ax, ay, sgn * mid, road)
        b.hidden = True
        obstacles.append(b)
    return Scenario(start=starts[start_pos], goal_point=goal, goal_area=goal_area,
                    allowed_goal_theta_difference=np.pi / 16, obstacles=obstacles)


# ---- motion primitives ---------------------------------------------------------------------------------------------
@dataclass
class MotionPrimitive:
    name: str
    total_length: float
    points: np.ndarray


def load_motion_primitives(version="bicycle_model") -> Dict[str, MotionPrimitive]:
    """Same dict (and iteration order: sorted file names) as main/lib/motion_primitive.py `load_motion_primitives`."""
    d = np.load(_DATA)
    names, pts, n, ln = d[version + "_names"], d[version + "_points"], d[version + "_n_points"], d[version + "_total_length"]
    return {str(nm): MotionPrimitive(str(nm), float(ln[i]), pts[i, :n[i]].copy()) for i, nm in enumerate(names)}


def resample_curve(points, dl, keep_last_point=True):
    """main/lib/trajectories.py:58-86 (scalar dl)."""
    d = points[1:, :2] - points[:-1, :2]
    step = np.append(0., np.sqrt(d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]))
    q = np.floor(step.cumsum() / dl).astype(int)
    mask = np.append(True, (q[1:] - q[:-1]) >= 1.)
    if keep_last_point:
        mask[-1] = True
    return points[mask].copy()


def collision_points(points, circle_centers, radius):
    """main/lib/motion_primitive_search_modified.py:35-52 for one primitive -> [n, 2]."""
    pts = resample_curve(points, radius)
    th = pts[:, 2]
    out = []
    for cx, cy in circle_centers:
        out.append(np.stack([np.cos(th) * cx - np.sin(th) * cy + pts[:, 0], np.sin(th) * cx + np.cos(th) * cy + pts[:, 1]], axis=1))
    return np.concatenate(out, axis=0)


class PrimitiveLibrary:
    """The primitives as the arrays the kernel takes (padded), built once."""

    def __init__(self, mps: Dict[str, MotionPrimitive], car_dimensions):
        self.names = list(mps)
        P = len(self.names)
        self.n_points = np.array([len(mps[n].points) for n in self.names], np.int32)
        self.NP = int(self.n_points.max())
        self.points = np.zeros((P, self.NP, 3))
        for i, n in enumerate(self.names):
            self.points[i, :self.n_points[i]] = mps[n].points
        self.length = np.array([float(mps[n].total_length) for n in self.names])
        cc = [collision_points(mps[n].points, np.asarray(car_dimensions.circle_centers), car_dimensions.radius) for n in self.names]
        self.NCC = max(len(c) for c in cc)
        # pad by repeating the last point: a duplicate point cannot change "any point inside any obstacle"
        self.cc = np.stack([np.concatenate([c, np.repeat(c[-1:], self.NCC - len(c), axis=0)]) for c in cc])


@dataclass
class PlanResult:
    status: np.ndarray          # [B] PLAN_*
    cost: np.ndarray            # [B]
    n_path: np.ndarray          # [B]
    mp_seq: np.ndarray          # [B, max_path] primitive index per edge, -1 padded
    nodes: np.ndarray           # [B, max_path, 3]
    n_expanded: np.ndarray      # [B]
    n_pushed: np.ndarray        # [B]
    traj: Optional[np.ndarray]  # [B, (max_path-1)*(NP-1), 3]
    n_traj: Optional[np.ndarray]

    def trajectory(self, b) -> np.ndarray:
        return self.traj[b, :self.n_traj[b]].copy()

    def path(self, b) -> List[Tuple[float, float, float]]:
        return [tuple(map(float, p)) for p in self.nodes[b, :self.n_path[b]]]


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def plan_batch(scenarios: Sequence[Scenario], library: PrimitiveLibrary, margin: float, device: int = 0,
               max_open: int = 65536, max_closed: int = 16384, max_path: int = 32, want_trajectory: bool = True,
               grow: int = 2) -> PlanResult:
    """A* for every scenario in one launch.  Scenarios that share their obstacle list (same list object) share one
    half-plane set on the device.  A search that runs out of open-list / closed-set capacity (PLAN_OVERFLOW) is run again,
    alone with the other overflowed ones, at 4x the capacities, up to `grow` times: the capacities are a first guess, not
    a limit of the planner (the reference's heap and dict grow as needed, main/lib/a_star.py:31-78)."""
    res = _plan_once(scenarios, library, margin, device, max_open, max_closed, max_path, want_trajectory)
    for _ in range(int(grow)):
        ov = np.flatnonzero(res.status == PLAN_OVERFLOW)
        if len(ov) == 0:
            break
        max_open, max_closed = 4 * max_open, 4 * max_closed
        sub = _plan_once([scenarios[b] for b in ov], library, margin, device, max_open, max_closed, max_path, want_trajectory)
        for name in ("status", "cost", "n_path", "mp_seq", "nodes", "n_expanded", "n_pushed", "traj", "n_traj"):
            if getattr(res, name) is not None:
                getattr(res, name)[ov] = getattr(sub, name)
    return res


def _plan_once(scenarios, library, margin, device, max_open, max_closed, max_path, want_trajectory) -> PlanResult:
    B = len(scenarios)
    sets, set_of, hp_list = {}, np.zeros(B, np.int32), []
    for b, sc in enumerate(scenarios):
        key = id(sc.obstacles)
        if key not in sets:
            sets[key] = len(hp_list)
            hp_list.append([o.to_convex(margin=margin) for o in sc.obstacles])
        set_of[b] = sets[key]
    S = len(hp_list)
    hp = np.zeros((S, MAX_OBS, MAX_HP, 3))
    n_hp = np.zeros((S, MAX_OBS), np.int32)
    n_obs = np.zeros(S, np.int32)
    for s, lst in enumerate(hp_list):
        if len(lst) > MAX_OBS:
            raise ValueError(f"at most {MAX_OBS} obstacles per scenario")
        n_obs[s] = len(lst)
        for o, h in enumerate(lst):
            hp[s, o, :len(h)] = h
            n_hp[s, o] = len(h)
    start = np.array([sc.start for sc in scenarios], float)
    goal = np.array([sc.goal_point for sc in scenarios], float)
    box = np.array([[*sc.goal_area.xy1, *sc.goal_area.xy2] for sc in scenarios], float)
    allowed = np.array([sc.allowed_goal_theta_difference for sc in scenarios], float)
    P, NP, NCC = len(library.names), library.NP, library.NCC
    res = PlanResult(status=np.zeros(B, np.int32), cost=np.zeros(B), n_path=np.zeros(B, np.int32),
                     mp_seq=np.zeros((B, max_path), np.int32), nodes=np.zeros((B, max_path, 3)),
                     n_expanded=np.zeros(B, np.int32), n_pushed=np.zeros(B, np.int32),
                     traj=np.zeros((B, (max_path - 1) * (NP - 1), 3)) if want_trajectory else None,
                     n_traj=np.zeros(B, np.int32) if want_trajectory else None)
    L = _lib.lib()
    rc = L.mpcb_plan_routes_host(
        int(device), B, _p(start), _p(goal), _p(box), _p(allowed), _p(set_of), S, _p(hp), _p(n_hp), _p(n_obs), P, NP,
        _p(np.ascontiguousarray(library.points)), _p(library.n_points), _p(library.length), NCC, _p(np.ascontiguousarray(library.cc)),
        int(max_open), int(max_closed), int(max_path), _p(res.status), _p(res.cost), _p(res.n_path), _p(res.mp_seq), _p(res.nodes),
        _p(res.n_expanded), _p(res.n_pushed), _p(res.traj), _p(res.n_traj))
    if rc != 0:
        raise _lib.MpcbError(f"mpcb_plan_routes_host failed (rc={rc}): {L.mpcb_plan_last_error().decode()}")
    return res


class MotionPrimitiveSearch:
    """main/lib/motion_primitive_search_modified.py:18-58 -- same constructor and `run()` -> (cost, path, trajectory)."""

    def __init__(self, scenario: Scenario, car_dimensions, mps: Dict[str, MotionPrimitive], margin: float, device: int = 0):
        self._scenario, self._margin, self._device = scenario, margin, device
        self._library = PrimitiveLibrary(mps, car_dimensions)
        self.result: Optional[PlanResult] = None

    def run(self, debug=False):
        r = plan_batch([self._scenario], self._library, self._margin, self._device)
        self.result = r
        if r.status[0] == PLAN_NO_SOLUTION:
            raise Exception("No solution found.")                        # main/lib/a_star.py:78
        if r.status[0] != PLAN_OK:
            raise _lib.MpcbError("planner capacity exceeded: raise max_open / max_closed / max_path in plan_batch")
        return float(r.cost[0]), r.path(0), r.trajectory(0)

    @property
    def primitive_names(self) -> List[str]:
        r = self.result
        return [self._library.names[k] for k in r.mp_seq[0, :r.n_path[0] - 1]]
