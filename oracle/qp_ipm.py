"""TEST INFRASTRUCTURE -- the float64 QP solver behind the oracle.

Dense Mehrotra predictor-corrector interior-point method for

    min 1/2 z'Pz + q'z   s.t.  Ez = e,  Gz <= h

followed by an active-set "polish" (an equality-constrained solve on the
identified active set, accepted only if it passes the full KKT certificate).
It stands in for cvxpy -> ECOS, which the reference calls at
main/lib/mpc.py:193-194 and which is neither vendored nor pinned nor
installable here: PARITY UNPINNED at the solver
boundary.  The QP is strictly convex in the inputs, so the optimum is unique;
this solver reaches it to ~1e-12, i.e. closer than ECOS's own 1e-8 tolerance.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
import this module.  It is the checker, never the product.
"""
import numpy as np

OPTIMAL = "optimal"
OPTIMAL_INACCURATE = "optimal_inaccurate"
INFEASIBLE = "infeasible"


def _kkt_solve(P, E, G, w, r1, r2):
    """Solve [[P + G'diag(w)G, E'], [E, 0]] [dz; dy] = [r1; r2]."""
    n = P.shape[0]
    p = E.shape[0]
    M = P + G.T @ (w[:, None] * G)
    if p == 0:
        return np.linalg.solve(M, r1), np.zeros(0)
    K = np.zeros((n + p, n + p))
    K[:n, :n] = M
    K[:n, n:] = E.T
    K[n:, :n] = E
    # tiny dual regularisation keeps the factorisation stable for redundant rows
    K[n:, n:] = -1e-13 * np.eye(p)
    sol = np.linalg.solve(K, np.concatenate([r1, r2]))
    return sol[:n], sol[n:]


def kkt_residuals(P, q, E, e, G, h, z, y, lam):
    """Return (stationarity, equality, inequality violation, complementarity, min dual)."""
    rd = P @ z + q + (E.T @ y if len(e) else 0.0) + G.T @ lam
    rp_e = E @ z - e if len(e) else np.zeros(0)
    slack = h - G @ z
    return (np.abs(rd).max(initial=0.0), np.abs(rp_e).max(initial=0.0),
            max(0.0, -slack.min(initial=0.0)),
            np.abs(slack * lam).max(initial=0.0), lam.min(initial=0.0))


def polish(P, q, E, e, G, h, z, y, lam, act_tol=1e-6):
    """Active-set refinement with certificate; returns (z, y, lam, ok)."""
    n = P.shape[0]
    slack = h - G @ z
    active = lam > slack  # strict complementarity guess
    Ga = G[active]
    ha = h[active]
    Ea = np.vstack([E, Ga]) if len(e) else Ga
    ea = np.concatenate([e, ha]) if len(e) else ha
    pa = Ea.shape[0]
    K = np.zeros((n + pa, n + pa))
    K[:n, :n] = P
    K[:n, n:] = Ea.T
    K[n:, :n] = Ea
    K[n:, n:] = -1e-14 * np.eye(pa)
    try:
        sol = np.linalg.solve(K, np.concatenate([-q, ea]))
    except np.linalg.LinAlgError:
        return z, y, lam, False
    z2 = sol[:n]
    mult = sol[n:]
    y2 = mult[:len(e)]
    lam2 = np.zeros_like(lam)
    lam2[active] = mult[len(e):]
    rd, rpe, viol, comp, lmin = kkt_residuals(P, q, E, e, G, h, z2, y2, lam2)
    ok = (rd < 1e-8 and rpe < 1e-9 and viol < 1e-9 and lmin > -1e-9)
    if ok:
        return z2, y2, lam2, True
    return z, y, lam, False


def solve_qp(P, q, E, e, G, h, tol=1e-11, max_iter=100, do_polish=True):
    """Returns dict(z, y, lam, s, status, iters, polished)."""
    P = np.asarray(P, float)
    q = np.asarray(q, float)
    n = P.shape[0]
    E = np.zeros((0, n)) if E is None else np.asarray(E, float).reshape(-1, n)
    e = np.zeros(0) if e is None else np.asarray(e, float).ravel()
    G = np.asarray(G, float).reshape(-1, n)
    h = np.asarray(h, float).ravel()
    m = G.shape[0]

    # Mehrotra-style starting point
    z, y = _kkt_solve(P, E, G, np.ones(m), -q + G.T @ h, e)
    s = h - G @ z
    lam = np.ones(m)
    shift = max(0.0, -1.5 * s.min(initial=1.0))
    s = s + shift
    s = np.maximum(s, 1e-2)

    status = OPTIMAL_INACCURATE
    it = 0
    scale = 1.0 + max(np.abs(q).max(initial=0.0), np.abs(h).max(initial=0.0))
    for it in range(1, max_iter + 1):
        rd = P @ z + q + G.T @ lam + (E.T @ y if len(e) else 0.0)
        rpe = (E @ z - e) if len(e) else np.zeros(0)
        rpi = G @ z + s - h
        mu = float(s @ lam) / m
        res = max(np.abs(rd).max(initial=0.0), np.abs(rpe).max(initial=0.0),
                  np.abs(rpi).max(initial=0.0))
        if res < tol * scale and mu < tol:
            status = OPTIMAL
            break
        w = lam / s

        def direction(rc):
            # rc = target for s*lam residual (vector)
            r1 = -rd + G.T @ (rc / s - w * rpi)
            dz, dy = _kkt_solve(P, E, G, w, r1, -rpe)
            ds = -rpi - G @ dz
            dlam = -(rc + lam * ds) / s
            return dz, dy, ds, dlam

        def max_step(v, dv):
            neg = dv < 0
            if not neg.any():
                return 1.0
            return min(1.0, float((-v[neg] / dv[neg]).min()))

        dz_a, dy_a, ds_a, dl_a = direction(s * lam)
        a_aff = min(max_step(s, ds_a), max_step(lam, dl_a))
        mu_aff = float((s + a_aff * ds_a) @ (lam + a_aff * dl_a)) / m
        sigma = (mu_aff / mu) ** 3 if mu > 0 else 0.0
        rc = s * lam + ds_a * dl_a - sigma * mu
        dz, dy, ds, dl = direction(rc)
        a = 0.995 * min(max_step(s, ds), max_step(lam, dl))
        a = min(a, 1.0)
        z = z + a * dz
        y = y + a * dy
        s = s + a * ds
        lam = lam + a * dl
        if not np.isfinite(z).all() or np.abs(z).max(initial=0.0) > 1e12:
            status = INFEASIBLE
            break
    else:
        # ran out of iterations
        rd = P @ z + q + G.T @ lam + (E.T @ y if len(e) else 0.0)
        rpi = G @ z + s - h
        if max(np.abs(rd).max(), np.abs(rpi).max()) > 1e-5 * scale:
            status = INFEASIBLE

    polished = False
    if status != INFEASIBLE and do_polish:
        z, y, lam, polished = polish(P, q, E, e, G, h, z, y, lam)
        if polished:
            s = h - G @ z
    return dict(z=z, y=y, lam=lam, s=s, status=status, iters=it, polished=polished)
