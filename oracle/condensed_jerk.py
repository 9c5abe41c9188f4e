"""TEST INFRASTRUCTURE -- numpy model of the CONDENSED form of main/lib/mpc_jerk.py that the CUDA jerk kernel uses.

main/lib/mpc_jerk.py differs from main/lib/mpc.py in the model (:61-86): a fifth state z with z_{t+1} = z_t + dt a_t
(`A[4,4] = 1`, `B[4,0] = dt`), a FREE initial value (only `x[:4, 0] == x0`, :193), and `A[2,4] = dt` on top of `B[2,0] = dt`:

    v_{t+1} = v_t + dt (a_t + z_t) =: v_t + dt e_t,          e = a + z,   z_k = z_0 + dt sum_{j<k} a_j

so the speed (and through it the position) sees the "effective" acceleration e = E [a; z_0] with
E[t, k] = [t == k] + dt [k < t], E[t, T] = 1.  Everything the tracking cost and the speed bounds need is therefore the
standard condensed model (oracle/condensed_model.py) evaluated at e, pulled back through E:

    H_jerk = E' H_track E + (input costs on a, incl. the jerk term dt^2 a_t^2 for t <= T-2, :190),   g likewise,

unknowns u = [a_0..a_{T-1} | delta_0..delta_{T-1} | z_0] (n = 2T + 1).  The constants are the ones that file hard-codes
(:16-39,167-171): w_perp 10, w_para 1, R (.01, .01), Rd (.3, 1), Q_v_yaw (0, .5), Qf (1, 1, 0, .5, 0) T, MAX_DECEL -5.
tests/test_oracle_golden.py checks this model against the goldens recorded from the unmodified module."""
import copy

import numpy as np

import condensed_model as cm
import mpc_oracle as mo
import qp_ipm


def jerk_cfg(T=13):
    cfg = mo.Cfg.from_dict({"T": T, "w_perp": 10.0, "w_para": 1.0, "R": [0.01, 0.01], "Rd": [0.3, 1.0], "Q_v_yaw": [0.0, 0.5],
                            "Qf": [1.0, 1.0, 0.0, 0.5], "MAX_DECEL": -5.0, "STOP_SPEED": 0.5 / 3.6})
    return cfg


class CondensedJerk:
    def __init__(self, xref, xbar, x0, reaches_end, cfg, L, dt, jerk_weight=1.0):
        self.T, self.dt, self.cfg, self.jw = cfg.T, dt, cfg, jerk_weight
        self.std = cm.Condensed(xref, xbar, x0, reaches_end, cfg, L, dt)
        tr = copy.copy(self.std)                       # tracking part only: no input costs
        tr.cfg = copy.copy(cfg)
        tr.cfg.Rd = (0.0, 0.0)
        tr.ra = np.zeros(self.T)
        tr.rd = np.zeros(self.T)
        self.track = tr
        T = self.T
        E = np.eye(T) + dt * np.tril(np.ones((T, T)), -1)
        self.E = np.hstack([E, np.ones((T, 1))])       # e = E [a; z0]

    def split(self, u):
        T = self.T
        return u[:T], u[T:2 * T], u[2 * T]

    def effective(self, a, z0):
        return self.E @ np.concatenate([a, [z0]])

    def hessian_gradient(self):
        """Dense H (n x n) and g (n) of J(u) = 1/2 u'Hu + g'u + const, u = [a | delta | z0]."""
        T, dt, cfg = self.T, self.dt, self.cfg
        A, B, C = self.track.hessian()                 # in (e, delta)
        g_e, g_d = self.track.grad(np.zeros(T), np.zeros(T))
        n = 2 * T + 1
        H = np.zeros((n, n))
        ia = list(range(T)) + [2 * T]                  # a and z0 both act through e
        idl = list(range(T, 2 * T))
        H[np.ix_(ia, ia)] = self.E.T @ A @ self.E
        H[np.ix_(idl, ia)] = B @ self.E
        H[np.ix_(ia, idl)] = (B @ self.E).T
        H[np.ix_(idl, idl)] = C
        g = np.zeros(n)
        g[ia] = self.E.T @ g_e
        g[idl] = g_d
        for blk0, r, rdw in ((0, self.std.ra, cfg.Rd[0]), (T, self.std.rd, cfg.Rd[1])):
            for i in range(T):
                H[blk0 + i, blk0 + i] += 2.0 * r[i]
                nb = (1 if i > 0 else 0) + (1 if i < T - 1 else 0)
                H[blk0 + i, blk0 + i] += 2.0 * rdw * nb
                if i > 0:
                    H[blk0 + i, blk0 + i - 1] -= 2.0 * rdw
                    H[blk0 + i - 1, blk0 + i] -= 2.0 * rdw
        for t in range(T - 1):                         # jerk: (z_{t+1} - z_t)^2 = dt^2 a_t^2, t = 0..T-2 (:190)
            H[t, t] += 2.0 * self.jw * dt * dt
        return H, g

    def constraints(self):
        """Dense G u <= h in the kernel's slot order (8 slots per stage, rows 4/5 missing at the last stage)."""
        T, dt, cfg = self.T, self.dt, self.cfg
        n = 2 * T + 1
        v0 = self.std.x0[2]
        rows, h = [], []

        def add(coef, rhs):
            rows.append(coef)
            h.append(rhs)
        cum = dt * np.tril(np.ones((T, T))) @ self.E    # v_{t+1} - v0 in terms of [a; z0]
        for t in range(T):
            ea = np.zeros(n); ea[t] = 1.0
            ed = np.zeros(n); ed[T + t] = 1.0
            add(ea, cfg.MAX_ACCEL); add(-ea, -cfg.MAX_DECEL)
            add(ed, mo.MAX_STEER); add(-ed, mo.MAX_STEER)
            if t < T - 1:
                dd = np.zeros(n); dd[T + t + 1] = 1.0; dd[T + t] = -1.0
                add(dd, cfg.MAX_DSTEER * dt); add(-dd, cfg.MAX_DSTEER * dt)
            cv = np.zeros(n); cv[:T] = cum[t, :T]; cv[2 * T] = cum[t, T]
            add(cv, mo.MAX_SPEED - v0); add(-cv, v0 - mo.MIN_SPEED)
        return np.array(rows), np.array(h)

    def solve(self):
        H, g = self.hessian_gradient()
        G, h = self.constraints()
        v0 = self.std.x0[2]
        if v0 > mo.MAX_SPEED + 1e-9 or v0 < mo.MIN_SPEED - 1e-9:
            return None
        r = qp_ipm.solve_qp(H, g, None, None, G, h)
        if r["status"] not in (qp_ipm.OPTIMAL, qp_ipm.OPTIMAL_INACCURATE):
            return None
        a, d, z0 = self.split(r["z"])
        x, y, v, ph = self.std.states(self.effective(a, z0), d)
        return dict(oa=a, od=d, z0=z0, ox=x, oy=y, ov=v, oyaw=ph, iters=r["iters"])
