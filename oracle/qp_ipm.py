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
VERBOSE = False


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


def _ipm_ineq(H, g, G, h, tol, max_iter):
    """Mehrotra predictor-corrector for min 1/2 w'Hw + g'w s.t. Gw <= h (H SPD), Cholesky normal equations.

    Runs until the scaled KKT merit max(|rd|/gscale, |rp|, mu) < tol or stops improving, and returns the
    best iterate seen; the caller's active-set polish then removes the remaining O(sqrt(mu)) error.
    """
    n = H.shape[0]
    m = G.shape[0]
    if m == 0:
        w = np.linalg.solve(H, -g)
        return w, np.zeros(0), np.zeros(0), OPTIMAL, 0
    # well-centred start: unit duals, slacks at least 1 (an infeasible-start method; rp absorbs the rest)
    w = np.linalg.solve(H + G.T @ G, -g + G.T @ h)
    s = np.maximum(h - G @ w, 1.0)
    lam = np.ones(m)
    gscale = 1.0 + np.abs(g).max(initial=0.0)
    best = None
    stall = 0
    it = 0
    for it in range(1, max_iter + 1):
        rd = H @ w + g + G.T @ lam
        rp = G @ w + s - h
        mu = float(s @ lam) / m
        merit = max(np.abs(rd).max() / gscale, np.abs(rp).max(), mu)
        if VERBOSE:
            print("ipm", it, np.abs(rd).max(), np.abs(rp).max(), mu)
        if best is None or merit < best[0]:
            best = (merit, w.copy(), s.copy(), lam.copy(), it)
            stall = 0
        else:
            stall += 1
        if merit < tol or stall >= 15 or mu < 1e-15:
            break
        ws = lam / s
        M = H + G.T @ (ws[:, None] * G)
        try:
            Lc = np.linalg.cholesky(M)
        except np.linalg.LinAlgError:
            break

        def direction(rc):
            rhs = -rd + G.T @ (rc / s - ws * rp)
            dw = np.linalg.solve(Lc.T, np.linalg.solve(Lc, rhs))
            res = rhs - M @ dw                      # one step of iterative refinement
            dw = dw + np.linalg.solve(Lc.T, np.linalg.solve(Lc, res))
            ds = -rp - G @ dw
            dl = -(rc + lam * ds) / s
            return dw, ds, dl

        def max_step(v, dv):
            neg = dv < 0
            return min(1.0, float((-v[neg] / dv[neg]).min())) if neg.any() else 1.0

        dw_a, ds_a, dl_a = direction(s * lam)
        a_aff = min(max_step(s, ds_a), max_step(lam, dl_a))
        mu_aff = float((s + a_aff * ds_a) @ (lam + a_aff * dl_a)) / m
        sigma = (mu_aff / mu) ** 3 if mu > 0 else 0.0
        dw, ds, dl = direction(s * lam + ds_a * dl_a - sigma * mu)
        a = min(1.0, 0.99 * min(max_step(s, ds), max_step(lam, dl)))
        w, s, lam = w + a * dw, s + a * ds, lam + a * dl
        if not np.isfinite(w).all() or np.abs(w).max(initial=0.0) > 1e12:
            return w, s, lam, INFEASIBLE, it
    merit, w, s, lam, _ = best
    if merit < 1e-8:
        status = OPTIMAL
    elif merit < 1e-5:
        status = OPTIMAL_INACCURATE
    else:
        status = INFEASIBLE
    return w, s, lam, status, it


def _polish_ineq(H, g, G, h, w, lam, max_rounds=60):
    """Active-set refinement on the reduced problem.

    Starts from the interior point's guess (lam > slack), solves the equality-constrained QP on that set and
    then repairs the set until the full KKT certificate holds: the most violated inactive row is added, the most
    negative multiplier is dropped (one change per round).  Dependent active rows are dropped by pivoted QR.
    Accepted only with the certificate, so a returned point is the optimum of the (strictly convex) QP whatever
    accuracy the interior point stopped at."""
    import scipy.linalg as sla
    n = H.shape[0]
    slack = h - G @ w
    active = set(np.where(lam > slack)[0].tolist())
    hscale = 1.0 + np.abs(h)
    for _ in range(max_rounds):
        act = np.array(sorted(active), dtype=int)
        if len(act):
            Ga = G[act]
            _, R, piv = sla.qr(Ga.T, mode="economic", pivoting=True)
            diag = np.abs(np.diag(R))
            rank = int((diag > 1e-9 * max(diag.max(initial=0.0), 1e-300)).sum())
            act = act[np.sort(piv[:rank])]
        Ga = G[act]
        k = len(act)
        K = np.zeros((n + k, n + k))
        K[:n, :n] = H
        K[:n, n:] = Ga.T
        K[n:, :n] = Ga
        rhs = np.concatenate([-g, h[act]])
        try:
            sol = np.linalg.solve(K, rhs)
            sol = sol + np.linalg.solve(K, rhs - K @ sol)
        except np.linalg.LinAlgError:
            return w, lam, False
        w2 = sol[:n]
        lam2 = np.zeros_like(lam)
        lam2[act] = sol[n:]
        viol = (G @ w2 - h) / hscale
        worst_v = int(np.argmax(viol)) if len(viol) else -1
        lam_floor = -1e-7 * (1.0 + np.abs(lam2).max(initial=0.0))
        worst_l = int(act[np.argmin(sol[n:])]) if k else -1
        if (worst_v < 0 or viol[worst_v] < 1e-9) and (k == 0 or lam2[worst_l] > lam_floor):
            return w2, lam2, True
        if worst_v >= 0 and viol[worst_v] >= 1e-9:
            active = set(act.tolist()) | {worst_v}
        else:
            active = set(act.tolist()) - {worst_l}
    return w, lam, False


def solve_qp(P, q, E, e, G, h, tol=1e-11, max_iter=120, do_polish=True):
    """Null-space elimination of Ez = e, then an inequality-only IPM on the reduced QP.

    Returns dict(z, y, lam, s, status, iters, polished); lam is for the rows of G (dropped
    constant rows get 0), y is not recovered (None).
    """
    import scipy.linalg as sla
    P = np.asarray(P, float)
    q = np.asarray(q, float)
    n = P.shape[0]
    E = np.zeros((0, n)) if E is None else np.asarray(E, float).reshape(-1, n)
    e = np.zeros(0) if e is None else np.asarray(e, float).ravel()
    G = np.asarray(G, float).reshape(-1, n)
    h = np.asarray(h, float).ravel()
    if E.shape[0]:
        Q, R = sla.qr(E.T, mode="full")               # E' = Q R
        p = E.shape[0]
        R1 = R[:p, :]
        zp = Q[:, :p] @ sla.solve_triangular(R1, e, trans="T")   # particular solution
        N = Q[:, p:]
    else:
        zp = np.zeros(n)
        N = np.eye(n)
    Hr = N.T @ P @ N
    Hr = 0.5 * (Hr + Hr.T)
    gr = N.T @ (P @ zp + q)
    Gr = G @ N
    hr = h - G @ zp
    # rows that do not depend on the free variables are constants: satisfied or infeasible
    rown = np.abs(Gr).max(axis=1) if Gr.size else np.zeros(0)
    const = rown < 1e-12
    if (hr[const] < -1e-9).any():
        return dict(z=zp, y=None, lam=np.zeros(len(h)), s=hr, status=INFEASIBLE, iters=0, polished=False)
    keep = ~const
    w, s, lam_k, status, it = _ipm_ineq(Hr, gr, Gr[keep], hr[keep], tol, max_iter)
    polished = False
    if status != INFEASIBLE and do_polish and keep.any():
        w, lam_k, polished = _polish_ineq(Hr, gr, Gr[keep], hr[keep], w, lam_k)
    lam = np.zeros(len(h))
    lam[keep] = lam_k
    z = zp + N @ w
    return dict(z=z, y=None, lam=lam, s=h - G @ z, status=status, iters=it, polished=polished)
