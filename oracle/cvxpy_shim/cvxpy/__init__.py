"""TEST INFRASTRUCTURE -- a stand-in for the `cvxpy` module.

The reference builds its MPC problem with cvxpy and solves it with ECOS
(main/lib/mpc.py:148-194).  Neither package is installed, pinned or
installable in this container, so this module
implements exactly the API subset those lines touch, assembles the QP

    min 1/2 z'Pz + q'z   s.t.  Ez = e,  Gz <= h

and hands it to oracle/qp_ipm.py.  With it on sys.path the reference's
main/lib/mpc.py runs UNMODIFIED, which is how tests/golden/ was generated.
PARITY UNPINNED: no real cvxpy/ECOS output exists to check against.

Not imported by the product; see oracle/README.md.
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.normpath(os.path.join(os.path.dirname(__file__), "..", "..")))
import qp_ipm  # noqa: E402

ECOS = "ECOS"
OPTIMAL = qp_ipm.OPTIMAL
OPTIMAL_INACCURATE = qp_ipm.OPTIMAL_INACCURATE
INFEASIBLE = qp_ipm.INFEASIBLE

_next_var_id = [0]

# last assembled problem, for tests that want to look at the matrices
LAST = {}


class Affine:
    """Vector-valued affine expression: sum_v coef[v] @ vec(v) + const, flattened row-major."""
    __array_ufunc__ = None
    __array_priority__ = 1000

    def __init__(self, coefs, const, shape):
        self.coefs = coefs          # dict: Variable -> (k, var.size) ndarray
        self.const = np.asarray(const, float).ravel()
        self.shape = tuple(shape)

    @property
    def size(self):
        return self.const.size

    # ---- helpers
    @staticmethod
    def _lift(other, like):
        if isinstance(other, Affine):
            return other
        arr = np.asarray(other, float)
        if arr.size == 1 and like.size != 1:
            arr = np.full(like.size, float(arr.ravel()[0]))
        return Affine({}, arr.ravel(), like.shape if arr.size == like.size else arr.shape)

    def _combine(self, other, sign):
        other = Affine._lift(other, self)
        a, b = self, other
        if a.size == 1 and b.size != 1:
            a = Affine({v: np.repeat(c, b.size, axis=0) for v, c in a.coefs.items()},
                       np.repeat(a.const, b.size), b.shape)
        if b.size == 1 and a.size != 1:
            b = Affine({v: np.repeat(c, a.size, axis=0) for v, c in b.coefs.items()},
                       np.repeat(b.const, a.size), a.shape)
        assert a.size == b.size, (a.shape, b.shape)
        coefs = {v: c.copy() for v, c in a.coefs.items()}
        for v, c in b.coefs.items():
            coefs[v] = coefs[v] + sign * c if v in coefs else sign * c
        return Affine(coefs, a.const + sign * b.const, a.shape)

    def __add__(self, o):
        return self._combine(o, 1.0)

    __radd__ = __add__

    def __sub__(self, o):
        return self._combine(o, -1.0)

    def __rsub__(self, o):
        return (-self)._combine(o, 1.0)

    def __neg__(self):
        return Affine({v: -c for v, c in self.coefs.items()}, -self.const, self.shape)

    def __mul__(self, k):
        k = float(k)
        return Affine({v: k * c for v, c in self.coefs.items()}, k * self.const, self.shape)

    __rmul__ = __mul__

    def __rmatmul__(self, M):
        M = np.asarray(M, float)
        assert len(self.shape) == 1 and M.ndim == 2 and M.shape[1] == self.size
        return Affine({v: M @ c for v, c in self.coefs.items()}, M @ self.const, (M.shape[0],))

    def __getitem__(self, key):
        idx = np.arange(self.size).reshape(self.shape)[key]
        shape = np.shape(idx)
        idx = np.ravel(idx)
        return Affine({v: c[idx] for v, c in self.coefs.items()}, self.const[idx], shape)

    # ---- constraints
    def __eq__(self, o):
        return Constraint(self - o, "eq")

    def __le__(self, o):
        return Constraint(self - o, "le")

    def __ge__(self, o):
        return Constraint(Affine._lift(o, self) - self, "le")

    __hash__ = None


class Variable(Affine):
    def __init__(self, shape):
        if isinstance(shape, int):
            shape = (shape,)
        self.id = _next_var_id[0]
        _next_var_id[0] += 1
        self.var_shape = tuple(shape)
        n = int(np.prod(shape))
        self.value = None
        Affine.__init__(self, {self: np.eye(n)}, np.zeros(n), shape)

    def __hash__(self):
        return self.id

    def __eq__(self, o):
        if isinstance(o, Variable):
            return self is o
        return Affine.__eq__(self, o)


class AbsExpr:
    def __init__(self, arg):
        self.arg = arg

    def __le__(self, c):
        return [Constraint(self.arg - c, "le"), Constraint(-self.arg - c, "le")]


class Constraint:
    def __init__(self, expr, kind):
        self.expr = expr     # expr == 0  or  expr <= 0
        self.kind = kind


class Quad:
    """Sum of quadratic forms (A z + b)' W (A z + b) plus a constant."""
    __array_ufunc__ = None

    def __init__(self, terms=(), const=0.0):
        self.terms = list(terms)
        self.const = const

    def __add__(self, o):
        if isinstance(o, Quad):
            return Quad(self.terms + o.terms, self.const + o.const)
        return Quad(self.terms, self.const + float(o))

    __radd__ = __add__

    def __mul__(self, k):                                   # scalar * expression (main/lib/mpc_jerk.py:190)
        k = float(k)
        assert k >= 0.0, "only non-negative scalings keep the objective convex"
        return Quad([(x, Pm * k) for x, Pm in self.terms], self.const * k)

    __rmul__ = __mul__


def quad_form(x, Pm):
    Pm = np.asarray(Pm, float)
    assert isinstance(x, Affine) and Pm.shape == (x.size, x.size)
    return Quad([(x, Pm)])


def square(x):
    assert isinstance(x, Affine) and x.size == 1
    return Quad([(x, np.ones((1, 1)))])


def abs(x):  # noqa: A001 - mirrors cvxpy.abs
    return AbsExpr(x)


class Minimize:
    def __init__(self, expr):
        self.expr = expr if isinstance(expr, Quad) else Quad([], float(expr))


class Problem:
    def __init__(self, objective, constraints=()):
        self.objective = objective
        flat = []
        for c in constraints:
            flat.extend(c if isinstance(c, list) else [c])
        self.constraints = flat
        self.status = None
        self.value = None
        self.solver_stats = None

    def _variables(self):
        vs = {}
        for x, _ in self.objective.expr.terms:
            for v in x.coefs:
                vs[v.id] = v
        for c in self.constraints:
            for v in c.expr.coefs:
                vs[v.id] = v
        return [vs[k] for k in sorted(vs)]

    def assemble(self):
        variables = self._variables()
        offs = {}
        n = 0
        for v in variables:
            offs[v] = n
            n += int(np.prod(v.var_shape))

        def dense(x):
            A = np.zeros((x.size, n))
            for v, c in x.coefs.items():
                A[:, offs[v]:offs[v] + c.shape[1]] += c
            return A, x.const

        P = np.zeros((n, n))
        q = np.zeros(n)
        c0 = self.objective.expr.const
        for x, W in self.objective.expr.terms:
            A, b = dense(x)
            Ws = 0.5 * (W + W.T)
            P += 2.0 * A.T @ Ws @ A
            q += 2.0 * A.T @ Ws @ b
            c0 += b @ Ws @ b
        Es, es, Gs, hs = [], [], [], []
        for c in self.constraints:
            A, b = dense(c.expr)
            if c.kind == "eq":
                Es.append(A)
                es.append(-b)
            else:
                Gs.append(A)
                hs.append(-b)
        E = np.vstack(Es) if Es else np.zeros((0, n))
        e = np.concatenate(es) if es else np.zeros(0)
        G = np.vstack(Gs) if Gs else np.zeros((0, n))
        h = np.concatenate(hs) if hs else np.zeros(0)
        return variables, offs, P, q, c0, E, e, G, h

    def solve(self, solver=None, verbose=False, **kw):
        variables, offs, P, q, c0, E, e, G, h = self.assemble()
        res = qp_ipm.solve_qp(P, q, E, e, G, h)
        self.status = res["status"]
        self.solver_stats = res
        LAST.clear()
        LAST.update(P=P, q=q, E=E, e=e, G=G, h=h, res=res)
        if self.status in (OPTIMAL, OPTIMAL_INACCURATE):
            z = res["z"]
            for v in variables:
                k = int(np.prod(v.var_shape))
                v.value = z[offs[v]:offs[v] + k].reshape(v.var_shape)
            self.value = 0.5 * z @ P @ z + q @ z + c0
        else:
            for v in variables:
                v.value = None
            self.value = None
        return self.value
