"""TEST INFRASTRUCTURE -- numpy model of the CONDENSED formulation the CUDA kernel uses.

Not the oracle (that is mpc_oracle.py, non-condensed, dense KKT).  This file is a
line-for-line executable description of the math in
mpc_for_av_at_intersection_b200/csrc/mpcb_kernels.cu: closed-form input->state
sensitivities of the linearised bicycle model (main/lib/mpc.py:58-79 with the steer
operating point fixed at 0), the Hessian from suffix-summed moments, the gradient by a
forward/adjoint scan, the structured constraint rows, and the Mehrotra iteration.
tests/test_condensed_model.py checks it against the oracle so that a kernel bug and a
derivation bug can be told apart.
"""
import numpy as np

import mpc_oracle as mo


class Condensed:
    def __init__(self, xref, xbar, x0, reaches_end, cfg, L, dt):
        T = cfg.T
        self.T, self.dt, self.cfg = T, dt, cfg
        self.x0 = np.asarray(x0, float)          # (x, y, v, yaw)
        vb, pb = xbar[2, :T], xbar[3, :T]         # operating point, stages 0..T-1
        self.c = dt * np.cos(pb)
        self.s = dt * np.sin(pb)
        self.p = dt * vb * np.sin(pb)
        self.q = dt * vb * np.cos(pb)
        self.b = dt * vb / L
        self.pb = pb
        # stage weights for states t = 1..T  (index t-1)
        re = np.asarray(reaches_end, bool)
        psi = xref[3, 1:]
        cs, sn = np.cos(psi), np.sin(psi)
        cp, sp = np.cos(psi + 0.5 * np.pi), np.sin(psi + 0.5 * np.pi)
        m11 = cfg.w_perp * cp * cp + cfg.w_para * cs * cs
        m12 = cfg.w_perp * cp * sp + cfg.w_para * cs * sn
        m22 = cfg.w_perp * sp * sp + cfg.w_para * sn * sn
        e = re[1:]
        self.m11 = np.where(e, cfg.Qf_scaled[0], m11)
        self.m12 = np.where(e, 0.0, m12)
        self.m22 = np.where(e, cfg.Qf_scaled[1], m22)
        self.wv = np.where(e, cfg.Qf_scaled[2], cfg.Q_v_yaw[0])
        self.wp = np.where(e, cfg.Qf_scaled[3], cfg.Q_v_yaw[1])
        self.ra = np.where(re[:T], 10.0, cfg.R[0])
        self.rd = np.where(re[:T], 10.0, cfg.R[1])
        self.xref = xref
        # prefix sums Z_t, W_t for t = 0..T  (Z_0 = 0)
        self.Zx = np.concatenate([[0.0], np.cumsum(self.c)])
        self.Zy = np.concatenate([[0.0], np.cumsum(self.s)])
        self.Wx = np.concatenate([[0.0], np.cumsum(-self.p)])
        self.Wy = np.concatenate([[0.0], np.cumsum(self.q)])

    # ---- linear dynamics and objective gradient by scans
    def states(self, a, d):
        T, dt = self.T, self.dt
        x0, y0, v0, p0 = self.x0
        v = v0 + dt * np.concatenate([[0.0], np.cumsum(a)])           # v_0..v_T
        ph = p0 + np.concatenate([[0.0], np.cumsum(self.b * d)])      # phi_0..phi_T
        dx = self.c * v[:T] - self.p * (ph[:T] - self.pb)
        dy = self.s * v[:T] + self.q * (ph[:T] - self.pb)
        x = x0 + np.concatenate([[0.0], np.cumsum(dx)])
        y = y0 + np.concatenate([[0.0], np.cumsum(dy)])
        return x, y, v, ph

    def grad(self, a, d):
        """dJ/da, dJ/ddelta with J the reference's cost (no 1/2)."""
        T, dt, cfg = self.T, self.dt, self.cfg
        x, y, v, ph = self.states(a, d)
        ex = x[1:] - self.xref[0, 1:]
        ey = y[1:] - self.xref[1, 1:]
        lx = 2.0 * (self.m11 * ex + self.m12 * ey)        # index t-1 for state t
        ly = 2.0 * (self.m12 * ex + self.m22 * ey)
        lv = 2.0 * self.wv * (v[1:] - self.xref[2, 1:])
        lp = 2.0 * self.wp * (ph[1:] - self.xref[3, 1:])
        # costates at state t = 1..T (index t-1): suffix sums
        px = np.cumsum(lx[::-1])[::-1]
        py = np.cumsum(ly[::-1])[::-1]
        # pv_t = lv_t + pv_{t+1} + c_t px_{t+1} + s_t py_{t+1};  c_t uses stage t (1..T-1)
        pxn = np.concatenate([px[1:], [0.0]])             # px_{t+1} for t = 1..T
        pyn = np.concatenate([py[1:], [0.0]])
        cv = np.concatenate([self.c[1:], [0.0]])          # c_t for t = 1..T (c_T unused)
        sv = np.concatenate([self.s[1:], [0.0]])
        pp = np.concatenate([self.p[1:], [0.0]])
        qq = np.concatenate([self.q[1:], [0.0]])
        pv = np.cumsum((lv + cv * pxn + sv * pyn)[::-1])[::-1]
        pph = np.cumsum((lp - pp * pxn + qq * pyn)[::-1])[::-1]
        ga = dt * pv                                       # dJ/da_k = dt * pv_{k+1}
        gd = self.b * pph
        ga = ga + 2.0 * self.ra * a
        gd = gd + 2.0 * self.rd * d
        da = np.diff(a)
        dd = np.diff(d)
        ga[:-1] -= 2.0 * cfg.Rd[0] * da
        ga[1:] += 2.0 * cfg.Rd[0] * da
        gd[:-1] -= 2.0 * cfg.Rd[1] * dd
        gd[1:] += 2.0 * cfg.Rd[1] * dd
        return ga, gd

    # ---- Hessian from suffix moments
    def hessian(self):
        T, dt, cfg = self.T, self.dt, self.cfg
        Zx, Zy, Wx, Wy = self.Zx[1:], self.Zy[1:], self.Wx[1:], self.Wy[1:]   # at states 1..T

        def suf(arr):   # S[m] = sum_{t=m..T} arr_t, m = 1..T+1 -> index m-1, S[T+1] = 0
            return np.concatenate([np.cumsum(arr[::-1])[::-1], [0.0]])
        m11, m12, m22 = self.m11, self.m12, self.m22
        S11, S12, S22 = suf(m11), suf(m12), suf(m22)
        MZx, MZy = m11 * Zx + m12 * Zy, m12 * Zx + m22 * Zy
        MWx, MWy = m11 * Wx + m12 * Wy, m12 * Wx + m22 * Wy
        S1zx, S1zy, S1wx, S1wy = suf(MZx), suf(MZy), suf(MWx), suf(MWy)
        S2zz, S2ww, S2wz = suf(Zx * MZx + Zy * MZy), suf(Wx * MWx + Wy * MWy), suf(Wx * MZx + Wy * MZy)
        Sv, Sp = suf(self.wv), suf(self.wp)
        A = np.zeros((T, T))
        B = np.zeros((T, T))
        C = np.zeros((T, T))
        for i in range(T):
            for j in range(T):
                m = max(i, j) + 1          # first state both inputs reach is m (>= k+1)
                k = m - 1                  # array index of S[m]
                zi = (self.Zx[i + 1], self.Zy[i + 1])
                zj = (self.Zx[j + 1], self.Zy[j + 1])
                wi = (self.Wx[i + 1], self.Wy[i + 1])
                wj = (self.Wx[j + 1], self.Wy[j + 1])

                def quad(S2, S1u, S1v, u, v):
                    # sum_t (U_t - u)' M_t (V_t - v)
                    return (S2[k] - (u[0] * S1v[0][k] + u[1] * S1v[1][k]) - (v[0] * S1u[0][k] + v[1] * S1u[1][k])
                            + u[0] * (S11[k] * v[0] + S12[k] * v[1]) + u[1] * (S12[k] * v[0] + S22[k] * v[1]))
                A[i, j] = 2.0 * dt * dt * (quad(S2zz, (S1zx, S1zy), (S1zx, S1zy), zi, zj) + Sv[k])
                C[i, j] = 2.0 * self.b[i] * self.b[j] * (quad(S2ww, (S1wx, S1wy), (S1wx, S1wy), wi, wj) + Sp[k])
                # row delta_i, column a_j:  U = W (with w_i), V = Z (with z_j)
                B[i, j] = 2.0 * self.b[i] * dt * quad(S2wz, (S1wx, S1wy), (S1zx, S1zy), wi, zj)
        for blk, r, rdw in ((A, self.ra, cfg.Rd[0]), (C, self.rd, cfg.Rd[1])):
            for i in range(T):
                blk[i, i] += 2.0 * r[i]
                nb = (1 if i > 0 else 0) + (1 if i < T - 1 else 0)
                blk[i, i] += 2.0 * rdw * nb
                if i > 0:
                    blk[i, i - 1] -= 2.0 * rdw
                    blk[i - 1, i] -= 2.0 * rdw
        return A, B, C

    # ---- constraints: slots per stage t: 0 a<=amax, 1 -a<=-amin, 2 d<=smax, 3 -d<=smax,
    #      4 d[t+1]-d[t]<=dmax, 5 -(..)<=dmax (t<=T-2), 6 v[t+1]<=vmax, 7 -v[t+1]<=-vmin
    def G_apply(self, a, d):
        T, dt = self.T, self.dt
        out = np.zeros((8, T))
        out[0], out[1], out[2], out[3] = a, -a, d, -d
        dd = np.concatenate([np.diff(d), [0.0]])
        out[4], out[5] = dd, -dd
        cs = dt * np.cumsum(a)
        out[6], out[7] = cs, -cs
        return out

    def h(self):
        T, cfg = self.T, self.cfg
        h = np.zeros((8, T))
        h[0], h[1] = cfg.MAX_ACCEL, -cfg.MAX_DECEL
        h[2] = h[3] = mo.MAX_STEER
        h[4] = h[5] = cfg.MAX_DSTEER * self.dt
        h[6] = mo.MAX_SPEED - self.x0[2]
        h[7] = self.x0[2] - mo.MIN_SPEED
        return h

    def mask(self):
        m = np.ones((8, self.T), bool)
        m[4:6, -1] = False
        return m

    def GT_apply(self, y):
        dt = self.dt
        ga = y[0] - y[1] + dt * np.cumsum((y[6] - y[7])[::-1])[::-1]
        r = y[4] - y[5]
        gd = y[2] - y[3] - r + np.concatenate([[0.0], r[:-1]])
        return ga, gd

    def GWG(self, w):
        T, dt = self.T, self.dt
        A = np.zeros((T, T))
        C = np.zeros((T, T))
        sv = np.cumsum((w[6] + w[7])[::-1])[::-1]
        for i in range(T):
            for j in range(T):
                A[i, j] = dt * dt * sv[max(i, j)]
        A += np.diag(w[0] + w[1])
        C += np.diag(w[2] + w[3])
        r = w[4] + w[5]
        for t in range(T - 1):
            C[t, t] += r[t]
            C[t + 1, t + 1] += r[t]
            C[t + 1, t] -= r[t]
            C[t, t + 1] -= r[t]
        return A, C

    # ---- Mehrotra predictor-corrector, the iteration the kernel runs
    def solve(self, a0=None, d0=None, tol_mu=1e-12, tol_p=1e-10, tol_d=1e-9, max_iter=40, verbose=False, eta_fn=None, init=None, warm=None,
              factor=None):
        """Stationarity test: |r_d| < tol_d + 1e-12 * (1 + |grad J(0)|_inf) (the gradient's rounding floor scales with it)."""
        T = self.T
        A, B, C = self.hessian()
        H = np.block([[A, B.T], [B, C]])
        msk = self.mask()
        h = self.h()
        a = np.zeros(T) if a0 is None else np.array(a0, float)
        d = np.zeros(T) if d0 is None else np.array(d0, float)
        g0a, g0d = self.grad(np.zeros(T), np.zeros(T))
        gscale = 1.0 + max(np.abs(g0a).max(), np.abs(g0d).max())
        # the kernel's start: slacks floored at 1, multipliers centred on them with mu0 = max(10, gscale / 100)
        s = np.where(msk, np.maximum(h - self.G_apply(a, d), 1.0), 1.0)
        lam = np.where(msk, max(10.0, 0.01 * gscale) / s, 0.0)
        if init == "legacy":                       # the round-1 start: lam = 1
            lam = np.where(msk, 1.0, 0.0)
            init = None
        if init is not None:                       # experiment hook: (slack floor, mu0) -> centred start lam = mu0 / s
            s = np.where(msk, np.maximum(h - self.G_apply(a, d), init[0]), 1.0)
            lam = np.where(msk, init[1] / s, 0.0)
        if warm is not None:                       # experiment hook: (lam_prev [8,T] or like, delta_s, delta_l)
            lp, ds_, dl_ = warm
            s = np.where(msk, np.maximum(h - self.G_apply(a, d), ds_), 1.0)
            lam = np.where(msk, np.maximum(lp, dl_), 0.0)
        m = msk.sum()
        status = 1
        mu_prev, eta_safe = 1e300, False
        for it in range(1, max_iter + 1):
            ga, gd = self.grad(a, d)
            gta, gtd = self.GT_apply(lam)
            rda, rdd = ga + gta, gd + gtd
            rp = np.where(msk, self.G_apply(a, d) + s - h, 0.0)
            mu = float((s * lam)[msk].sum()) / m
            res_d = max(np.abs(rda).max(), np.abs(rdd).max())
            res_p = np.abs(rp).max()
            if verbose:
                print(it, res_d, res_p, mu)
            if res_d < tol_d + 1e-12 * gscale and res_p < tol_p and mu < tol_mu:
                status = 0
                it -= 1
                break
            w = np.where(msk, lam / s, 0.0)
            WA, WC = self.GWG(w)
            M = H + np.block([[WA, np.zeros((T, T))], [np.zeros((T, T)), WC]])
            try:
                if factor is None:
                    Lc = np.linalg.cholesky(M)
                    solveM = lambda rhs: np.linalg.solve(Lc.T, np.linalg.solve(Lc, rhs))
                else:
                    # experiment hook (scripts/fp32_factor_experiment.py): ("f32", k) = symmetric Jacobi scaling, Cholesky and
                    # triangular solves in float32, k steps of iterative refinement against the float64 matrix
                    _, nref = factor
                    dsc = 1.0 / np.sqrt(np.diag(M))
                    Ms = (M * dsc[:, None]) * dsc[None, :]
                    L32 = np.linalg.cholesky(Ms.astype(np.float32))
                    if not np.isfinite(L32).all():
                        raise np.linalg.LinAlgError("float32 factor")

                    def solve32(r):
                        y = np.linalg.solve(L32, r.astype(np.float32))
                        return np.linalg.solve(L32.T, y).astype(np.float64)

                    def solveM(rhs, Ms=Ms, dsc=dsc):
                        b = rhs * dsc
                        x = solve32(b)
                        for _ in range(nref):
                            x = x + solve32(b - Ms @ x)
                        return x * dsc
            except np.linalg.LinAlgError:
                status = 0 if (res_d < 1e-7 * gscale and res_p < 1e-8 and mu < 1e-8 * max(1.0, 1e-4 * gscale)) else 2      # gap relative to the cost scale, as in the kernels
                it -= 1
                break

            def direction(rc):
                y = np.where(msk, rc / s - w * rp, 0.0)
                ya, yd = self.GT_apply(y)
                rhs = np.concatenate([-rda + ya, -rdd + yd])
                du = solveM(rhs)
                da_, dd_ = du[:T], du[T:]
                ds = np.where(msk, -rp - self.G_apply(da_, dd_), 0.0)
                dl = np.where(msk, -(rc + lam * ds) / s, 0.0)
                return da_, dd_, ds, dl

            def maxstep(v, dv):
                neg = (dv < 0) & msk
                return min(1.0, float((-v[neg] / dv[neg]).min())) if neg.any() else 1.0

            da_, dd_, ds, dl = direction(s * lam)
            al = min(maxstep(s, ds), maxstep(lam, dl))
            mu_aff = float(((s + al * ds) * (lam + al * dl))[msk].sum()) / m
            sigma = (mu_aff / mu) ** 3
            da_, dd_, ds, dl = direction(s * lam + ds * dl - sigma * mu)
            # the kernel's adaptive fraction to the boundary (mpcb_step.cuh): 0.99 -> 1 as the gap closes, floored so that one
            # step cannot push the gap far below its target; back to 0.99 for good once an aggressive step failed to halve the gap
            if eta_fn is None:
                if mu_prev < 0.01 and mu > 0.5 * mu_prev:
                    eta_safe = True
                mu_prev = mu
                eta = 0.99 if eta_safe else 1.0 - min(0.01, max(mu, 0.1 * tol_mu / mu))
            else:
                eta = eta_fn(mu)
            al = min(1.0, eta * min(maxstep(s, ds), maxstep(lam, dl)))
            a, d, s, lam = a + al * da_, d + al * dd_, s + al * ds, lam + al * dl
        x, y, v, ph = self.states(a, d)
        return dict(oa=a, od=d, ox=x, oy=y, ov=v, oyaw=ph, iters=it, status=status, s=s, lam=lam)
