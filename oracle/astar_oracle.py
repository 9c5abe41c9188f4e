"""TEST INFRASTRUCTURE -- restatement of the reference's motion-primitive A* planner.

Follows main/lib/a_star.py:31-78 (the search loop, heap order, lazy deletion, predecessor dictionary) and
main/lib/motion_primitive_search_modified.py:101-149 (neighbour expansion: collision-check the primitive's circle-centre
points against every obstacle's half-planes, then transform the primitive's last point), :64-86 (goal test and heuristic),
:151-165 (trajectory assembly), main/lib/obstacles.py:157-176 (`check_collision`), :108-117 (box distance),
main/lib/linalg.py:4-55 (transform), main/lib/maths.py:4-10 (`normalize_angle`), main/lib/trajectories.py:58-86
(`resample_curve` for the collision points) and :11-55 (circle-centre offsets).

One deliberate difference: the reference multiplies points by the transform with `@` (OpenBLAS, whose kernels fuse
multiply-adds differently from machine to machine); here every product and sum is written out in a fixed order
(`(px*c + py*(-s)) + x`), the order the CUDA kernel uses too.  Pinned against tests/golden/astar_intersection.npz
(84 searches recorded from the unmodified reference): primitive index sequences, costs and expansion counts must be
identical, nodes and trajectories equal to 1e-9.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module."""
import math
from heapq import heappop, heappush

import numpy as np


# sin / cos with a FIXED sequence of IEEE double operations (no libm, no FMA): Cody-Waite reduction by pi/2 in three
# parts and the classic degree-13 / degree-14 minimax polynomials on [-pi/4, pi/4].  The CUDA kernel
# (csrc/mpcb_astar.cu: sincos_fixed) performs the same operations in the same order, so node coordinates -- and with them
# every tie the search breaks -- are bit-identical on both sides.  Accuracy ~1 ulp; the reference's own np.cos / BLAS
# products differ from either by as much.
_PIO2_1, _PIO2_2, _PIO2_3 = 1.57079632673412561417e+00, 6.07710050630396597660e-11, 2.02226624871116645580e-21
_INVPIO2 = 6.36619772367581382433e-01
_S = (-1.66666666666666324348e-01, 8.33333333332248946124e-03, -1.98412698298579493134e-04, 2.75573137070700676789e-06,
      -2.50507602534068634195e-08, 1.58969099521155010221e-10)
_C = (4.16666666666666019037e-02, -1.38888888888741095749e-03, 2.48015872894767294178e-05, -2.75573143513906633035e-07,
      2.08757232129817482790e-09, -1.13596475577881948265e-11)


def sincos_fixed(x):
    k = float(round(x * _INVPIO2))                       # round half to even = rint
    r = ((x - k * _PIO2_1) - k * _PIO2_2) - k * _PIO2_3
    z = r * r
    ps = _S[0] + z * (_S[1] + z * (_S[2] + z * (_S[3] + z * (_S[4] + z * _S[5]))))
    sn = r + (r * z) * ps
    pc = _C[0] + z * (_C[1] + z * (_C[2] + z * (_C[3] + z * (_C[4] + z * _C[5]))))
    cs = (1.0 - 0.5 * z) + (z * z) * pc
    q = int(k) & 3
    if q == 0:
        return sn, cs
    if q == 1:
        return cs, -sn
    if q == 2:
        return -sn, -cs
    return -cs, sn


def normalize_angle(theta):
    """main/lib/maths.py:4-10."""
    theta = theta % math.tau
    if theta >= math.pi:
        theta -= math.tau
    return theta


def resample_points(points, dl):
    """main/lib/trajectories.py:58-86 with keep_last_point=True."""
    d = points[1:, :2] - points[:-1, :2]
    step = np.append(0., np.sqrt(d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]))
    q = np.floor(step.cumsum() / dl).astype(int)
    mask = np.append(True, (q[1:] - q[:-1]) >= 1.)
    mask[-1] = True
    return points[mask].copy()


def collision_points(points, circle_centers, radius):
    """motion_primitive_search_modified.py:35-52: resampled primitive -> circle-centre trajectories, concatenated
    (main/lib/trajectories.py:11-55).  Returns [n, 2] object-space points."""
    pts = resample_points(points, radius)
    th = pts[:, 2]
    out = []
    for cx, cy in circle_centers:
        ox = np.cos(th) * cx - np.sin(th) * cy
        oy = np.sin(th) * cx + np.cos(th) * cy
        out.append(np.stack([ox + pts[:, 0], oy + pts[:, 1]], axis=1))
    return np.concatenate(out, axis=0)


class AStarOracle:
    def __init__(self, start, goal, goal_box, allowed, halfplanes, mp_points, mp_n_points, mp_length, cc_points, blas=False):
        """halfplanes: list of [m, 3] arrays (a, b, c: inside where a x + b y + c <= 0 for every row);
        mp_points [P, NPmax, 3], mp_length [P], cc_points [P, NCC, 2].
        blas=True multiplies with `@` exactly as the reference does (main/lib/linalg.py:41-44, obstacles.py:172) and
        squares with `**`: on the machine that recorded the golden file this reproduces even the expansion counts, which
        shows the search logic is restated exactly; blas=False is the fixed-order arithmetic the CUDA kernel is checked
        against (same primitive sequences and costs; expansion counts may differ by a few ties)."""
        self.blas = blas
        self.start = tuple(float(v) for v in start)
        self.gx, self.gy, self.gth = (float(v) for v in goal)
        self.box = tuple(float(v) for v in goal_box)
        self.allowed = float(allowed)
        self.hp = [np.asarray(h, float) for h in halfplanes]
        self.mp_points, self.mp_n, self.mp_len, self.cc = mp_points, mp_n_points, [float(v) for v in mp_length], cc_points
        self.edge_mp = {}
        self.n_expanded = 0

    # -- motion_primitive_search_modified.py:64-86
    def is_goal(self, node):
        x, y, th = node
        x1, y1, x2, y2 = self.box
        dx = max(x1 - x, 0, x - x2)
        dy = max(y1 - y, 0, y - y2)
        return math.sqrt(dx * dx + dy * dy) <= 1e-5 and abs(th - self.gth) <= self.allowed

    def heuristic(self, node):
        x, y, th = node
        if self.blas:
            dxy = float(np.sqrt((x - self.gx) ** 2 + (y - self.gy) ** 2))
        else:
            dxy = math.sqrt((x - self.gx) * (x - self.gx) + (y - self.gy) * (y - self.gy))
        dth = min(abs(th - self.gth), abs(th - self.gth) - self.allowed / 2)
        return dxy + 2.7 * dth

    @staticmethod
    def _transform(node, px, py):
        x, y, th = node
        s, c = sincos_fixed(th)
        return (px * c + py * (-s)) + x, (px * s + py * c) + y

    def neighbors(self, node):
        """:101-149."""
        x, y, th = node
        s, c = sincos_fixed(th)
        mtx = np.array([[np.cos(th), -np.sin(th), x], [np.sin(th), np.cos(th), y], [0, 0, 1]])      # linalg.py:21-23
        for m in range(len(self.mp_len)):
            cc = self.cc[m]
            if self.blas:
                w = (np.append(cc[:, :2], np.ones((len(cc), 1)), axis=1) @ mtx.T)[:, :2]
                wx, wy = w[:, 0], w[:, 1]
                pts1 = np.vstack([w.T, np.ones((len(cc),))])
            else:
                wx = (cc[:, 0] * c + cc[:, 1] * (-s)) + x
                wy = (cc[:, 0] * s + cc[:, 1] * c) + y
            collides = False
            for h in self.hp:                                   # obstacles.py:157-176
                if self.blas:
                    val = h @ pts1
                else:
                    val = (h[:, 0:1] * wx[None, :] + h[:, 1:2] * wy[None, :]) + h[:, 2:3]
                if bool(np.any(np.all(val <= 0, axis=0))):
                    collides = True
                    break
            if collides:
                continue
            last = self.mp_points[m, self.mp_n[m] - 1]
            if self.blas:
                nx, ny = (np.append(np.atleast_2d(last)[:, :2], np.ones((1, 1)), axis=1) @ mtx.T)[0, :2].tolist()
            else:
                nx, ny = self._transform(node, float(last[0]), float(last[1]))
            nb = (nx, ny, normalize_angle(float(last[2]) + th))
            self.edge_mp[node, nb] = m
            yield self.mp_len[m], nb

    def run(self, max_expanded=200000):
        """main/lib/a_star.py:31-78.  Returns (cost, path nodes, primitive index sequence) or None ("No solution found.")."""
        start = self.start
        q = [(0, 0, start, start)]
        pred = {}
        while q:
            gh, g, node, p = heappop(q)
            if node in pred and g >= pred[node][0]:
                continue
            self.n_expanded += 1
            if self.n_expanded > max_expanded:
                raise RuntimeError("expansion limit")
            pred[node] = g, p
            if self.is_goal(node):
                path = [node]
                while node != start:
                    path.append(p)
                    node, p = p, pred[p][1]
                path.reverse()
                return g, path, [self.edge_mp[a, b] for a, b in zip(path[:-1], path[1:])]
            for cost, nb in self.neighbors(node):
                ng = g + cost
                if nb not in pred or ng < pred[nb][0]:
                    heappush(q, (ng + self.heuristic(nb), ng, nb, node))
        return None

    def trajectory(self, path, seq):
        """:151-165: every primitive's points but the last, transformed to its start node."""
        out = []
        for node, m in zip(path[:-1], seq):
            pts = self.mp_points[m, :self.mp_n[m] - 1]
            x, y, th = node
            s, c = sincos_fixed(th)
            out.append(np.stack([(pts[:, 0] * c + pts[:, 1] * (-s)) + x, (pts[:, 0] * s + pts[:, 1] * c) + y, pts[:, 2] + th], axis=1))
        return np.concatenate(out, axis=0) if out else np.zeros((0, 3))
