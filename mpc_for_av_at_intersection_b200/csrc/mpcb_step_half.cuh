// mpcb_step_half.cuh -- the fused MPC step for SHORT horizons (T + 1 <= 16): TWO problems per warp, one per half-warp.
//
// Same reference rows, same math, same outputs as mpcb_step.cuh (A2..A8: main/lib/mpc.py:86-239,280-326,
// main/lib/trajectories.py:100-126, main/lib/simulation.py:35-47).  At the reference's default horizon T = 13 the
// lane-per-stage kernel leaves 19 of 32 lanes idle in every instruction; here lane l = lane & 15 of half h = lane >> 4 is
// stage l of problem h, every shuffle / scan / reduction runs with width 16, each half owns a WarpSmem<T>, and the
// rank-T update on the FP64 tensor cores (a full-warp instruction) is issued twice, once per half's tiles.
//
// The two problems of a warp run in LOCKSTEP: a diverged half could not wait in __syncwarp, so nothing here branches on
// per-problem data around a warp-level primitive.  A half without a live problem (no problem for the slot, finished, goal
// reached) mirrors its partner's problem with every global store suppressed; a problem whose nearest-index lookup hits the
// reference's Exception keeps computing on its old index and only reports the status; the interior-point loop runs until
// BOTH halves have stopped; a stopped half's iterate, multipliers and bookkeeping are frozen (every update is a select on
// !stop) while it idles through the instructions -- its step directions may go NaN, nobody reads them.
#pragma once
#include "mpcb_step.cuh"

namespace mpcb {

#define MPCB_HW 16
__device__ __forceinline__ double shfl_h(double v, int src) { return __shfl_sync(MPCB_FULL, v, src, MPCB_HW); }
__device__ __forceinline__ double shfl_up1_h(double v, int l) {
    const double n = __shfl_up_sync(MPCB_FULL, v, 1, MPCB_HW);
    return l >= 1 ? n : 0.0;
}
__device__ __forceinline__ double shfl_dn1_h(double v, int l) {
    const double n = __shfl_down_sync(MPCB_FULL, v, 1, MPCB_HW);
    return l < MPCB_HW - 1 ? n : 0.0;
}
__device__ __forceinline__ double scan_up_h(double v, int l) {          // inclusive prefix sum over lanes 0..l of the half
#pragma unroll
    for (int d = 1; d < MPCB_HW; d <<= 1) {
        const double n = __shfl_up_sync(MPCB_FULL, v, d, MPCB_HW);
        if (l >= d) v += n;
    }
    return v;
}
__device__ __forceinline__ double scan_dn_h(double v, int l) {          // inclusive suffix sum over lanes l..15 of the half
#pragma unroll
    for (int d = 1; d < MPCB_HW; d <<= 1) {
        const double n = __shfl_down_sync(MPCB_FULL, v, d, MPCB_HW);
        if (l + d < MPCB_HW) v += n;
    }
    return v;
}
__device__ __forceinline__ double grp_sum(double v) {
#pragma unroll
    for (int d = MPCB_HW / 2; d >= 1; d >>= 1) v += __shfl_xor_sync(MPCB_FULL, v, d);
    return v;
}
__device__ __forceinline__ double grp_max(double v) {
#pragma unroll
    for (int d = MPCB_HW / 2; d >= 1; d >>= 1) v = fmax(v, __shfl_xor_sync(MPCB_FULL, v, d));
    return v;
}
__device__ __forceinline__ double grp_max_nn(double v) {              // no NaN among the inputs (see dmax_run)
#pragma unroll
    for (int d = MPCB_HW / 2; d >= 1; d >>= 1) v = dmax_run(v, __shfl_xor_sync(MPCB_FULL, v, d));
    return v;
}
__device__ __forceinline__ bool grp_any(bool v, int half) {             // any lane of THIS half
    return ((__ballot_sync(MPCB_FULL, v) >> (16 * half)) & 0xffffu) != 0;
}

// [phase: nearest idx]
// main/lib/trajectories.py:100-126, forward=True, on one half-warp (see nearest_index_in_direction); executed by both
// halves together: the loop bound is the longer of the two scans
__device__ __forceinline__ int nearest_index_h(const double* __restrict__ cx, const double* __restrict__ cy, int start, int nc,
                                               double x, double y, int l) {
    const int n = nc - start;
    double bd[3] = {1e300, 1e300, 1e300};
    int bi[3] = {0x7fffffff, 0x7fffffff, 0x7fffffff};
    int nmax = n;
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) nmax = max(nmax, __shfl_xor_sync(MPCB_FULL, nmax, o));
    for (int j0 = 0; j0 < nmax; j0 += MPCB_HW) {
        const int j = start + j0 + l;
        if (j < nc) {
            const double dx = __dsub_rn(cx[j], x), dy = __dsub_rn(cy[j], y);
            top3_insert(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), j, bd, bi);
        }
    }
    int ind[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        double d = bd[0];
        int i = bi[0];
#pragma unroll
        for (int o = MPCB_HW / 2; o >= 1; o >>= 1) {
            const double d2 = __shfl_xor_sync(MPCB_FULL, d, o);
            const int i2 = __shfl_xor_sync(MPCB_FULL, i, o);
            if (d2 < d || (d2 == d && i2 < i)) { d = d2; i = i2; }
        }
        ind[r] = i;
        if (bi[0] == i) {
            bd[0] = bd[1]; bi[0] = bi[1];
            bd[1] = bd[2]; bi[1] = bi[2];
            bd[2] = 1e300; bi[2] = 0x7fffffff;
        }
    }
    if (n <= 1) return start;
    if (n == 2) return start + 1;
    if (abs(ind[1] - ind[2]) == 2) return ind[0];
    if (abs(ind[0] - ind[1]) == 1) return max(ind[0], ind[1]);
    return -1;
}

// [phase: G ops]
template <int T>
__device__ __forceinline__ void G_apply_h(double a, double d, double dt, int l, double (&g)[8]) {
    const double dn = shfl_dn1_h(d, l);
    const double dd = (l < T - 1) ? dn - d : 0.0;
    const double sc = scan_up_h(a, l);                                // (a warp primitive: never inside the select)
    const double cs = l < T ? dt * sc : 0.0;                          // every row is zero on the idle lanes (a = d = 0 there)
    g[0] = a; g[1] = -a; g[2] = d; g[3] = -d; g[4] = dd; g[5] = -dd; g[6] = cs; g[7] = -cs;
}
template <int T>
__device__ __forceinline__ void GT_apply_h(const double (&y)[8], double dt, int l, double& ga, double& gd) {
    ga = y[0] - y[1] + dt * scan_dn_h(y[6] - y[7], l);
    const double r = y[4] - y[5];
    gd = y[2] - y[3] - r + shfl_up1_h(r, l);
}

// [phase: kernel head/inputs]
// One warp, two problems.  b_in: this half's problem or -1 (no problem for the slot).  Returns the interior-point iterations
// this half's problem took (0 if it was skipped).
template <int T, int WARPS, bool VAR>
__device__ __forceinline__ int mpc_step_half_body(const StepParams& p, const int b_in) {
    static_assert(T >= 3 && T + 1 <= MPCB_HW, "two problems per warp need T + 1 <= 16");
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int l = lane & (MPCB_HW - 1), half = lane >> 4;
    WarpSmem<T>* const smw = reinterpret_cast<WarpSmem<T>*>(smem_raw) + 2 * wid;      // the two halves' blocks
    WarpSmem<T>& sm = smw[half];
    const bool act = l < T;
    const double dt = p.dt;

    // ---- which problem this half works on, and whether its results count ---------------------------------------
    bool live = b_in >= 0 && !(p.done && p.done[b_in] != 0);
    if (live && p.fuse_goal) {       // MPC.is_goal (main/lib/mpc.py:310-326) as the loop head of mpc_intersection.py:97
        const double gd = hypot(p.state[4 * b_in] - p.goal[2 * b_in], p.state[4 * b_in + 1] - p.goal[2 * b_in + 1]);
        const bool isgoal = gd <= p.goal_dis && abs(p.target_ind[b_in] - p.ncourse[b_in]) < 5 && fabs(p.state[4 * b_in + 3]) <= p.stop_speed;
        if (isgoal) { if (l == 0) p.done[b_in] = 1; live = false; }
    }
    if (!__any_sync(MPCB_FULL, live)) return 0;
    const int b_partner = __shfl_xor_sync(MPCB_FULL, live ? b_in : -1, 16);
    const int b = live ? b_in : b_partner;               // a half without a live problem mirrors its partner's (stores suppressed)
    const double x0 = p.state[4 * b + 0], y0 = p.state[4 * b + 1], yaw0 = p.state[4 * b + 2], v0 = p.state[4 * b + 3];
    const int route = p.route[b];
    const int nc = p.ncourse[b];
    const double L = p.L ? p.L[b] : p.wheelbase;
    const double dl = p.dl[route];
    const double* __restrict__ cx = p.courses + (size_t)route * 3 * p.ncmax;
    const double* __restrict__ cy = cx + p.ncmax;
    const double* __restrict__ cyaw = cy + p.ncmax;
    int tind = p.target_ind[b];
    const int warm = p.warm[b];
    double oa = (warm && act) ? p.oa[(size_t)b * T + l] : 0.0;     // mpc.py:222-224
    double od = (warm && act) ? p.od[(size_t)b * T + l] : 0.0;

    // cost weights: the handle's, or this problem's own row (VAR kernels only)
    double w_perp = p.w_perp, w_para = p.w_para, R0 = p.R0, R1 = p.R1, Rd0 = p.Rd0, Rd1 = p.Rd1, Qv = p.Qv, Qyaw = p.Qyaw,
           Qf0 = p.Qf0, Qf1 = p.Qf1, Qf2 = p.Qf2, Qf3 = p.Qf3;
    int vcut = 0x7fffffff;
    if constexpr (VAR) {
        if (p.pw) {
            const double* __restrict__ w = p.pw + 12 * (size_t)b;
            w_perp = w[0]; w_para = w[1]; R0 = w[2]; R1 = w[3]; Rd0 = w[4]; Rd1 = w[5]; Qv = w[6]; Qyaw = w[7];
            Qf0 = w[8]; Qf1 = w[9]; Qf2 = w[10]; Qf3 = w[11];
        }
        if (p.vcut) vcut = p.vcut[b];
    }
    double xr_v = 0.0;                                 // xref[2, l]: speed reference (mpc_with_speed.py:104), else 0
    int status = MPCB_ST_OPTIMAL, iters = 0;
    bool solved = false, have_ov = false, nearest_fail = false, failed = false;
    double sx = 0.0, sy = 0.0, sv = 0.0, sph = 0.0;    // solution states at index l (0..T)
    double xr_x = 0.0, xr_y = 0.0, xr_yaw = 0.0;       // xref at state index l (0..T)

    for (int pass = 0; pass < p.max_iter; ++pass) {
        // [phase: A2]
        // ---- A2: reference lookup -----------------------------------------------------------------
        const int ni = nearest_index_h(cx, cy, tind, nc, x0, y0, l);
        if (ni < 0) nearest_fail = true;                  // the reference raises (trajectories.py:120): reported at the end,
        else if (!nearest_fail) tind = ni;                // meanwhile this half keeps its old index and idles along
        double trav = 0.0;
        {   // |ov| * dt with ov = max(v0, 10/3.6) on the first pass (mpc.py:96-98), the previous pass's speeds afterwards;
            // one loop for both cases (a shuffle of a warp-uniform value returns it): np.cumsum is sequential
            const double term = __dmul_rn(fabs(have_ov ? sv : fmax(v0, p.min_ref_speed)), dt);
            double acc = 0.0;
#pragma unroll 1
            for (int i = 0; i <= T; ++i) {
                const double ti = shfl_h(term, i);
                acc = (i == 0) ? ti : __dadd_rn(acc, ti);
                if (i == l) trav = acc;
            }
        }
        int idx = (int)rint(__ddiv_rn(trav, dl)) + tind;    // np.rint: half to even
        idx = min(idx, nc - 1);
        if (l > T) idx = nc - 1;
        idx = max(idx, 0);
        xr_x = cx[idx]; xr_y = cy[idx]; xr_yaw = cyaw[idx];
        if constexpr (VAR) xr_v = idx < vcut ? p.ref_speed : 0.0;
        const bool reach = (idx == nc - 1);

        // [phase: A3]
        // ---- A3: rollout of the previous (unshifted) solution: operating point (vbar, phibar) -------
        const double dcl = clampd(od, -p.max_steer, p.max_steer);
        const double tnd = act ? tan(dcl) : 0.0;
        double vb = v0, vcur = v0;
#pragma unroll 1
        for (int t = 0; t < T - 1; ++t) {
            const double at = shfl_h(oa, t);
            vcur = clampd(__dadd_rn(vcur, __dmul_rn(at, dt)), p.min_speed, p.max_speed);
            if (l == t + 1) vb = vcur;
        }
        const double yterm = __dmul_rn(__dmul_rn(__ddiv_rn(vb, L), tnd), dt);
        double pb = yaw0, pcur = yaw0;
#pragma unroll 1
        for (int t = 0; t < T - 1; ++t) {
            pcur = __dadd_rn(pcur, shfl_h(yterm, t));
            if (l == t + 1) pb = pcur;
        }

        // [phase: A4/A5]
        // ---- A4: linearisation coefficients of stage `l` -----------------------------------------
        double sn, cs;
        sincos(pb, &sn, &cs);
        const double dtv = dt * vb;
        const double c_ = act ? dt * cs : 0.0, s_ = act ? dt * sn : 0.0;
        const double pp = act ? dtv * sn : 0.0, qq = act ? dtv * cs : 0.0;
        const double bb = act ? dtv / L : 0.0;
        if (!act) pb = 0.0;
        const double cn = shfl_dn1_h(c_, l), snn = shfl_dn1_h(s_, l), ppn = shfl_dn1_h(pp, l), qqn = shfl_dn1_h(qq, l);

        // ---- A5: stage weights for state l+1, input weights for stage l -------------------------
        const double rx = shfl_dn1_h(xr_x, l), ry = shfl_dn1_h(xr_y, l), rpsi = shfl_dn1_h(xr_yaw, l);
        double rv = 0.0;
        if constexpr (VAR) rv = shfl_dn1_h(xr_v, l);
        const bool reach_n = __shfl_down_sync(MPCB_FULL, (int)reach, 1, MPCB_HW) != 0;
        double m11 = 0.0, m12 = 0.0, m22 = 0.0, wv = 0.0, wp = 0.0;
        if (act) {
            if (!reach_n) {
                double s1, c1, s2, c2;
                sincos(rpsi + 0.5 * 3.141592653589793, &s1, &c1);     // mpc.py:160
                sincos(rpsi, &s2, &c2);                                // mpc.py:164
                m11 = w_perp * (c1 * c1) + w_para * (c2 * c2);
                m12 = w_perp * (c1 * s1) + w_para * (c2 * s2);
                m22 = w_perp * (s1 * s1) + w_para * (s2 * s2);
                wv = Qv; wp = Qyaw;
            } else { m11 = Qf0; m22 = Qf1; wv = Qf2; wp = Qf3; }
        }
        const double ra_w = act ? (reach ? p.r_end : R0) : 0.0;     // mpc.py:177-180
        const double rd_w = act ? (reach ? p.r_end : R1) : 0.0;

        // [phase: H build]
        // ---- condensed Hessian from suffix moments ----------------------------------------------------
        const double Zx = scan_up_h(c_, l), Zy = scan_up_h(s_, l);
        const double Wx = scan_up_h(-pp, l), Wy = scan_up_h(qq, l);
        {
            const double MZx = m11 * Zx + m12 * Zy, MZy = m12 * Zx + m22 * Zy;
            const double MWx = m11 * Wx + m12 * Wy, MWy = m12 * Wx + m22 * Wy;
            const double S11 = scan_dn_h(m11, l), S12 = scan_dn_h(m12, l), S22 = scan_dn_h(m22, l);
            const double S1zx = scan_dn_h(MZx, l), S1zy = scan_dn_h(MZy, l);
            const double S1wx = scan_dn_h(MWx, l), S1wy = scan_dn_h(MWy, l);
            const double S2zz = scan_dn_h(Zx * MZx + Zy * MZy, l);
            const double S2ww = scan_dn_h(Wx * MWx + Wy * MWy, l);
            const double S2wz = scan_dn_h(Wx * MZx + Wy * MZy, l);
            const double Sv = scan_dn_h(wv, l), Sp = scan_dn_h(wp, l);
            const double gAx = S11 * Zx + S12 * Zy - S1zx, gAy = S12 * Zx + S22 * Zy - S1zy;
            const double gCx = S11 * Wx + S12 * Wy - S1wx, gCy = S12 * Wx + S22 * Wy - S1wy;
            const double hA = S2zz - (Zx * S1zx + Zy * S1zy) + Sv;
            const double hC = S2ww - (Wx * S1wx + Wy * S1wy) + Sp;
            const double hB = S2wz - (Wx * S1zx + Wy * S1zy);
            const double hh = S2wz - (Zx * S1wx + Zy * S1wy);
            __syncwarp();
            if (act) {
                sm.bc[0][l] = Zx; sm.bc[1][l] = Zy; sm.bc[2][l] = Wx; sm.bc[3][l] = Wy;
                sm.bc[4][l] = bb; sm.bc[5][l] = hh; sm.bc[6][l] = gAx; sm.bc[7][l] = gAy;
            }
            __syncwarp();
            const double kA = 2.0 * dt * dt, kB = 2.0 * bb * dt, kC = 2.0 * bb;
            const int nb = (l > 0 ? 1 : 0) + (l < T - 1 ? 1 : 0);
#pragma unroll 1
            for (int j = 0; j < T; ++j) {
                const double zxj = sm.bc[0][j], zyj = sm.bc[1][j], wxj = sm.bc[2][j], wyj = sm.bc[3][j];
                double Aij = kA * (hA + zxj * gAx + zyj * gAy);
                double Cij = kC * sm.bc[4][j] * (hC + wxj * gCx + wyj * gCy);
                const double Blo = hB + zxj * gCx + zyj * gCy;
                const double Bup = sm.bc[5][j] + Wx * sm.bc[6][j] + Wy * sm.bc[7][j];
                const double Bij = kB * ((j <= l) ? Blo : Bup);
                if (j == l) { Aij += 2.0 * ra_w + 2.0 * Rd0 * nb; Cij += 2.0 * rd_w + 2.0 * Rd1 * nb; }
                if (j == l - 1) { Aij -= 2.0 * Rd0; Cij -= 2.0 * Rd1; }
                if (act) {
                    sm.HB[WarpSmem<T>::colB(j) + l] = Bij;
                    if (j <= l) {                             // symmetric blocks: both triangles are stored
                        sm.HA[WarpSmem<T>::colB(j) + l] = Aij; sm.HC[WarpSmem<T>::colB(j) + l] = Cij;
                        sm.HA[WarpSmem<T>::colB(l) + j] = Aij; sm.HC[WarpSmem<T>::colB(l) + j] = Cij;
                    }
                }
            }
            __syncwarp();
        }

        if (act) {
            const double tv[6] = {c_, s_, pp, qq, bb, pb};
#pragma unroll
            for (int i = 0; i < 6; ++i) sm.tab[i][l] = tv[i];
        }
        __syncwarp();
        auto tld = [&](int i) { return act ? sm.tab[i][l] : 0.0; };
        // [phase: predict/grad defs]
        // ---- linear prediction and objective gradient by scans ------------------------------------------
        // returns the states at index l+1 (lanes < T); see oracle/condensed_model.py states()/grad()
        auto predict = [&](double a, double d, double& xs, double& ys, double& vs, double& phs) {
            const double c_ = tld(0), s_ = tld(1), pp = tld(2), qq = tld(3), bb = tld(4), pb = tld(5);
            const double sa = scan_up_h(a, l);
            const double bd_ = bb * d;
            const double sp_ = scan_up_h(bd_, l);
            vs = v0 + dt * sa;
            phs = yaw0 + sp_;
            const double vt = v0 + dt * (sa - a);
            const double dph = (yaw0 + (sp_ - bd_)) - pb;
            xs = x0 + scan_up_h(c_ * vt - pp * dph, l);
            ys = y0 + scan_up_h(s_ * vt + qq * dph, l);
        };
        // gradient of the cost by the forward/adjoint scans: used ONCE, at u = 0 (g0); inside the interior-point
        // loop the gradient is g0 + H u from the stored Hessian (a 12-instruction rolled loop instead of 10 scans)
        auto gradient_scan = [&](double a, double d, double& ga, double& gd) {
            double xs, ys, vs, phs;
            predict(a, d, xs, ys, vs, phs);
            const double ex = xs - rx, ey = ys - ry;
            const double lx = 2.0 * (m11 * ex + m12 * ey), ly = 2.0 * (m12 * ex + m22 * ey);
            // xref[2,:] = 0 in main/lib/mpc.py:102-105 (rv == 0); cv[idx] in mpc_with_speed.py:104
            const double lv = 2.0 * wv * (VAR ? vs - rv : vs), lp = 2.0 * wp * (phs - rpsi);
            const double px = scan_dn_h(lx, l), py = scan_dn_h(ly, l);
            const double pxn = shfl_dn1_h(px, l), pyn = shfl_dn1_h(py, l);
            const double pv = scan_dn_h(lv + cn * pxn + snn * pyn, l);
            const double pph = scan_dn_h(lp - ppn * pxn + qqn * pyn, l);
            const double ap = shfl_up1_h(a, l), an = shfl_dn1_h(a, l);
            const double dp = shfl_up1_h(d, l), dn = shfl_dn1_h(d, l);
            const double fa = (l > 0 ? a - ap : 0.0) - (l < T - 1 ? an - a : 0.0);
            const double fd = (l > 0 ? d - dp : 0.0) - (l < T - 1 ? dn - d : 0.0);
            ga = act ? dt * pv + 2.0 * ra_w * a + 2.0 * Rd0 * fa : 0.0;
            gd = act ? bb * pph + 2.0 * rd_w * d + 2.0 * Rd1 * fd : 0.0;
        };

        // [phase: ipm init + grad matvec]
        // ---- A6: interior point on  min J(u)  s.t.  G u <= h ----------------------------------------------
        double hh8[8];
        hh8[0] = p.max_accel; hh8[1] = -p.max_decel; hh8[2] = p.max_steer; hh8[3] = p.max_steer;
        hh8[4] = p.dsteer_dt; hh8[5] = p.dsteer_dt; hh8[6] = p.max_speed - v0; hh8[7] = v0 - p.min_speed;
        bool mk[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) mk[c] = (c == 4 || c == 5) ? (l < T - 1) : act;
        const double m_cnt = 8.0 * T - 2.0;
        // x[2,0] <= MAX_SPEED and >= MIN_SPEED are constant rows (mpc.py:186-188): infeasible if violated
        bool feasible0 = (hh8[6] >= -1e-9) && (hh8[7] >= -1e-9);

        double a = oa, d = od;
        double gscale, g0a, g0d;
        gradient_scan(0.0, 0.0, g0a, g0d);
        gscale = 1.0 + grp_max(fmax(fabs(g0a), fabs(g0d)));
        // Starting point: slacks s = max(h - G u, 1) at the previous solution, multipliers CENTRED on them, lam = mu0 / s,
        // with the gap mu0 = max(10, gscale / 100) tied to the size of the cost gradient.  A vehicle far off its path has
        // gradients of 1e6..1e7 and multipliers of that order at the optimum: from lam = 1 the iteration needs 60-70 tiny
        // steps to grow them (the failures of tests/test_gpu_bench_workloads.py at T = 80), from a scaled start ~20; for
        // ordinary problems (gscale 1e2..1e4) the count is unchanged (numpy twin: 9.24 -> 9.09 per solve on the bench loop).
        double s[8], lam[8];
        {
            double g[8];
            G_apply_h<T>(a, d, dt, l, g);
            const double mu0 = fmax(10.0, 0.01 * gscale);
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                s[c] = mk[c] ? fmax(hh8[c] - g[c], 1.0) : 1.0;
                lam[c] = mk[c] ? mu0 * rcp(s[c]) : 0.0;
            }
            // Resident sessions: the multipliers of the previous control step's solve, floored at 1 (the next QP is the same
            // problem one stage on, so the active rows keep their multipliers; everything else restarts from 1).  The
            // optimum is unique, the answer does not depend on the start -- only the iteration count does: 9.15 -> 8.13
            // per solve on the bench loop (numpy twin, profiles/r02_NOTES.md).
            if (p.lam_ws && warm && pass == 0) {
                const double* __restrict__ lw = p.lam_ws + (size_t)b * 8 * T;
                const double fl = fmax(1.0, 1e-5 * gscale);
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    const double lp = act ? lw[c * T + l] : 0.0;
                    lam[c] = mk[c] ? fmax(lp, fl) : 0.0;
                }
            }
        }
        // gradient inside the interior-point loop: g0 + H u from the stored cost Hessian (column j of every block is
        // read down the lanes; the transposed cross block along row `l`)
        const int grow = act ? l : 0;                     // idle lanes (T <= l < 16) read row 0: valid memory, finite values
        auto gradient = [&](double a, double d, double& ga, double& gd) {
            using SMg = WarpSmem<T>;
            double sa = g0a, sd = g0d;
            const double* const hbr = &sm.HB[SMg::colB(grow)];                   // hbr[j] = B(steer j, accel l)
#pragma unroll 4
            for (int j = 0; j < T; ++j) {
                const double aj = shfl_h(a, j), dj = shfl_h(d, j);
                const int cj = SMg::colB(j) + grow;
                sa = fma(sm.HA[cj], aj, sa); sa = fma(hbr[j], dj, sa);
                sd = fma(sm.HB[cj], aj, sd); sd = fma(sm.HC[cj], dj, sd);
            }
            ga = act ? sa : 0.0; gd = act ? sd : 0.0;
        };
        double res_d = 1e300, res_p = 1e300, mu = 1e300;
        bool converged = false, broke = false;
        int it = 0;
        double mu_prev = 1e300;
        bool eta_safe = false;
        // Residual bookkeeping.  A Newton step of length al scales BOTH linear residuals by (1 - al): r_p exactly (the
        // slack step is defined from it), r_d up to the backward error of the LDL' solve.  The loop therefore carries
        // the residuals forward by that factor and re-evaluates them from scratch (Hessian mat-vec, G and G' scans)
        // only every sixteenth iteration (every fourth cost 1.8 % for nothing: the verification refresh before acceptance is what guarantees accuracy) and -- always -- before it accepts convergence or gives up, so the reported
        // accuracy is that of freshly computed residuals.
        double rda = 0.0, rdd = 0.0, rp[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) rp[c] = 0.0;
        bool verified = false;
        // Best VERIFIED iterate so far (freshly computed residuals), kept in the (now idle) assembly tables: if the loop ends
        // without meeting the configured tolerances -- iteration cap, pivot breakdown of the unpivoted LDL' at extreme lam/s
        // ratios, a step poisoned by such a factorisation -- the answer is that iterate and ITS residual class, the way a
        // CPU interior point keeps its best point.  merit < 1: the 1e-8 class the reference's ECOS calls optimal;
        // < 100: ECOS's reduced (OPTIMAL_INACCURATE) class, which main/lib/mpc.py:196 accepts too.
        // (the merit lives in shared memory too: the kernel sits at the 255-register limit and these values are touched
        // only where the residuals are fresh -- the first iteration, every sixteenth, and the end game)
        double& best_m = sm.bc[5][0];                   // warp-uniform: every l writes the same value
        best_m = 1e300;
        int since = 0;                                  // iterations since the carried dual residual fell under its tolerance
        // Two nested loops (residual evaluation outside the hot loop, see mpcb_step.cuh), both with WARP-UNIFORM control: the
        // outer one runs while either half is still iterating, the inner one is left as soon as either half needs fresh
        // residuals (the other half gets an extra evaluation, which is harmless).  `stop` freezes a half: from then on its
        // bookkeeping (iteration count, status flags, best iterate, committed answer) is not touched any more.
        // [phase: residuals (outer loop)]
        bool stop = !feasible0 || failed || nearest_fail;
        while (__any_sync(MPCB_FULL, !stop)) {
            {
                double ga, gd, gta, gtd, g[8];
                gradient(a, d, ga, gd);
                GT_apply_h<T>(lam, dt, l, gta, gtd);
                G_apply_h<T>(a, d, dt, l, g);
                double rpm = 0.0, sl = 0.0;
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    const double r = mk[c] ? g[c] + s[c] - hh8[c] : 0.0;
                    if (!stop) rp[c] = r;
                    rpm = fmax(rpm, fabs(r));
                    sl += s[c] * lam[c];
                }
                const double rd_new = grp_max(fmax(fabs(ga + gta), fabs(gd + gtd))), rp_new = grp_max(rpm);
                const double m = fmax(fmax(rd_new * (1e8 * rcp(gscale)), rp_new * 1e9), grp_sum(sl) / m_cnt * (1e9 * rcp(fmax(1.0, 1e-4 * gscale))));
                const double bm = best_m;
                __syncwarp();                            // every lane has read the old merit before any lane replaces it
                if (!stop) {
                    rda = ga + gta; rdd = gd + gtd; res_d = rd_new; res_p = rp_new;
                    verified = true;
                    if (m < bm && m < 100.0) {
                        best_m = m;
                        if (act) { sm.bc[6][l] = a; sm.bc[7][l] = d; }
                    }
                }
            }
        // [phase: gap + stopping tests]
        for (;;) {
            {
                double sl = 0.0;
#pragma unroll
                for (int c = 0; c < 8; ++c) sl += s[c] * lam[c];
                const double mu_new = grp_sum(sl) / m_cnt;
                if (!stop) mu = mu_new;
            }
            bool leave = false;                          // this half wants fresh residuals before it decides
            if (!stop) {
                if (res_d < p.tol_d + 1e-12 * gscale && res_p < p.tol_p && mu < p.tol_mu) {
                    if (verified) { converged = true; stop = true; }
                    else leave = true;
                } else if (broke || it >= p.ipm_max_iter || !(mu == mu) || mu < 1e-15) {
                    if (verified) stop = true;
                    else leave = true;
                }
            }
            if (__any_sync(MPCB_FULL, leave) || !__any_sync(MPCB_FULL, !stop)) break;

            // [phase: W, rhs]
            // ---- M = H + G' W G into registers (rows `l` of A, B, C) ---------------------------------
            double sinv[8], w[8];
#pragma unroll
            for (int c = 0; c < 8; ++c) { sinv[c] = rcp(s[c]); w[c] = lam[c] * sinv[c]; }
            using SM = WarpSmem<T>;
            double dC, wr_prev, wv_suf, dA;
            {
                wv_suf = dt * dt * scan_dn_h(w[6] + w[7], l);                // dt^2 * sum_{t>=l}
                const double wr = w[4] + w[5];
                wr_prev = shfl_up1_h(wr, l);
                dA = w[0] + w[1];
                dC = w[2] + w[3] + wr + wr_prev;
            }
            // affine right-hand side: -r_d + G'(lam - W r_p)
            double ra, rdl;
            {
                double y[8], ya, yd;
#pragma unroll
                for (int c = 0; c < 8; ++c) y[c] = lam[c] - w[c] * rp[c];
                GT_apply_h<T>(y, dt, l, ya, yd);
                ra = ya - rda; rdl = yd - rdd;
            }
            // [phase: factor driver+syrk]
            // ---- LDL' of M = H + G'WG, forward solve of the predictor right-hand side fused -------------------
            bool bad = false;
            double inva = 0.0, invd = 0.0;
            {
                double R[T], Q[T];                               // accel-block row and cross-block row of this l
#pragma unroll
                for (int j = 0; j < T; ++j) {                    // idle lanes read valid memory, their rows are never used
                    R[j] = sm.HA[SM::colB(j) + l] + wv_suf;   // dt^2 L'WL: entry (i, j <= i) is this l's suffix sum
                    Q[j] = sm.HB[SM::colB(j) + l];
                }
                double a0 = R[0] + (l == 0 ? dA : 0.0);
                double ivm = rcp(a0);
                __syncwarp();
                // three window sizes (measured: 2 segments 8.02 M solves/s, 3: 8.14, 4: 8.12, 6: 8.10 at B = 4096 -- wasted FMAs on
                // dead columns against bytes of hot-loop code)
                constexpr int K1 = T / 3, K2 = (2 * T) / 3;
                ldl_accel_steps<T, T - 1>(sm, R, Q, 0, K1, l, wv_suf, dA, ra, rdl, bad, inva, a0, ivm);
                ldl_accel_steps<T, T - 1 - K1>(sm, R, Q, K1, K2, l, wv_suf, dA, ra, rdl, bad, inva, a0, ivm);
                ldl_accel_steps<T, T - 1 - K2>(sm, R, Q, K2, T, l, wv_suf, dA, ra, rdl, bad, inva, a0, ivm);
            }
            {
                // steer block: assembled only now (the accel rows are dead), rank-T update C -= L21 D^-1 L21',
                // then its own LDL'
                // The rank-T update runs on the FP64 tensor cores: S = -L21 D^-1 L21' as 8x8 tiles of
                // mma.m8n8k4 (lower tiles only), operands gathered straight from the stored cross factor -- no operand
                // broadcast through the LSU, 256 FMAs per instruction.  The tiles go through a scratch square (the
                // steer factor's storage, dead until the LDL' below) to reach the l-per-row layout.
                __syncwarp();
                {
                    // a full-warp instruction: issued once per half's problem, all 32 lanes gathering that problem's operands
                    constexpr int NTL = (T + 7) / 8;                       // 8-row tiles
                    constexpr int KB = (T + 3) / 4;                        // 4-pivot blocks
                    const int fr = lane >> 2, fk = lane & 3;               // fragment row / k index of this lane
#pragma unroll 1
                    for (int hh = 0; hh < 2; ++hh) {
                        WarpSmem<T>& sh = smw[hh];
                        double acc[NTL * (NTL + 1) / 2][2];
#pragma unroll
                        for (int t = 0; t < NTL * (NTL + 1) / 2; ++t) { acc[t][0] = 0.0; acc[t][1] = 0.0; }
#pragma unroll 1
                        for (int kb = 0; kb < KB; ++kb) {
                            const int k = 4 * kb + fk;
                            const bool kv = (T % 4 == 0) || k < T;
                            const int kc = kv ? k : 0;
                            const double* col = &sh.LB[SM::colB(kc) + fr];
                            const double ivk = sh.iz[kc].x;
                            double fa[NTL], fb[NTL];
#pragma unroll
                            for (int x = 0; x < NTL; ++x) {
                                const double raw = col[8 * x];             // rows >= T: pad / next column, dead entries only
                                fb[x] = kv ? raw : 0.0;
                                fa[x] = kv ? -raw * ivk : 0.0;
                            }
#pragma unroll
                            for (int I = 0; I < NTL; ++I)
#pragma unroll
                                for (int J = 0; J <= I; ++J) dmma_m8n8k4(acc[I * (I + 1) / 2 + J][0], acc[I * (I + 1) / 2 + J][1], fa[I], fb[J]);
                        }
                        double* S = sh.LC;                                 // [row][SB], rows and columns < T
#pragma unroll
                        for (int I = 0; I < NTL; ++I)
#pragma unroll
                            for (int J = 0; J <= I; ++J) {
                                const int r = 8 * I + fr, c = 8 * J + 2 * fk;
                                if (r < T && c + 1 < T + (T & 1)) *reinterpret_cast<double2*>(&S[r * SM::SB + c]) =
                                    make_double2(acc[I * (I + 1) / 2 + J][0], acc[I * (I + 1) / 2 + J][1]);
                            }
                    }
                }
                __syncwarp();
                // G'WG of the steer block (diagonal dC, sub-diagonal -wr_prev) goes into this l's row of the scratch, i.e.
                // into the window before the elimination starts instead of onto the pivot-to-pivot chain of every step
                if (act) {
                    double* Sw = &sm.LC[l * SM::SB];
                    Sw[l] += dC;
                    if (l >= 1) Sw[l - 1] -= wr_prev;
                }
                double C[T];
                {
                    const double* Sr = &sm.LC[(act ? l : 0) * SM::SB];
#pragma unroll
                    for (int j = 0; j + 1 < T; j += 2) {                  // entries above the diagonal are never used
                        const double2 sv = *reinterpret_cast<const double2*>(Sr + j);
                        C[j] = sm.HC[SM::colB(j) + l] + sv.x;
                        C[j + 1] = sm.HC[SM::colB(j + 1) + l] + sv.y;
                    }
                    if (T & 1) C[T - 1] = sm.HC[SM::colB(T - 1) + l] + Sr[T - 1];
                }
                __syncwarp();                                              // the scratch is overwritten by the factor below
                double c0 = C[0];
                double ivm = rcp(c0);
                constexpr int K1 = T / 3, K2 = (2 * T) / 3;
                ldl_steer_steps<T, T - 1>(sm, C, 0, K1, l, act, dC, wr_prev, rdl, bad, invd, c0, ivm);
                ldl_steer_steps<T, T - 1 - K1>(sm, C, K1, K2, l, act, dC, wr_prev, rdl, bad, invd, c0, ivm);
                ldl_steer_steps<T, T - 1 - K2>(sm, C, K2, T, l, act, dC, wr_prev, rdl, bad, invd, c0, ivm);
            }
            __syncwarp();
            // every l's own pivot reciprocals come back from the published {1/d, z} pairs (one load each instead of a
            // select per elimination step); a pivot that was not positive and finite shows in its reciprocal
            inva = act ? sm.iz[l].x : 0.0;
            invd = act ? sm.iz[T + l].x : 0.0;
            bad = bad || (act && !(inva > 0.0 && inva < 1e300 && invd > 0.0 && invd < 1e300));
            {   // a breakdown in THIS half's factorisation: no step from it -- verify the iterate it has, then stop
                bool leave2 = false;
                if (grp_any(bad, half) && !stop) { broke = true; if (verified) stop = true; else leave2 = true; }
                if (__any_sync(MPCB_FULL, leave2) || !__any_sync(MPCB_FULL, !stop)) break;
            }
            if (!act) { ra = 0.0; rdl = 0.0; }
            // [phase: solves]
            // Triangular solves with the pivots in rolled loops.  The factor blocks are zero on and above the diagonal,
            // so every step is shuffle -> FMA with no predicate; the running right-hand sides are kept PRE-SCALED by
            // 1/d_i (the coefficient is scaled instead, off the dependency chain), which makes the value a l
            // broadcasts at its own step the finished unknown.
            const int lrow = act ? l : 0;                 // idle lanes read column 0 (finite entries; their own results are zeroed by inva = invd = 0)
            const double* const LAr = &sm.LA[SM::colL(lrow)];        // LAr[k] = (row k, column l)
            const double* const LCr = &sm.LC[SM::colL(lrow)];
            const double* const LBr = &sm.LB[SM::colB(lrow)];        // LBr[k] = (steer row k, accel column l)
            auto back_solve = [&](double ha, double hd, double& xa, double& xd) {      // L' x = h, h = D^-1 z
#pragma unroll 4
                for (int k = T - 1; k >= 0; --k) {
                    const double xk = shfl_h(hd, k);
                    const double uc = LCr[k] * invd, ub = LBr[k] * inva;
                    hd = fma(-uc, xk, hd);
                    ha = fma(-ub, xk, ha);
                }
                xd = hd;
#pragma unroll 4
                for (int k = T - 1; k >= 1; --k) {
                    const double xk = shfl_h(ha, k);
                    const double ua = LAr[k] * inva;
                    ha = fma(-ua, xk, ha);
                }
                xa = ha;
            };
            auto fwd_solve = [&](double& fa, double& fd) {     // L z = r, factor read from shared memory; returns D^-1-free z
                const double* colA = &sm.LA[lrow];
                const double* colB = &sm.LB[lrow];
                double ga = fa * inva;
#pragma unroll 4
                for (int k = 0; k < T; ++k) {
                    const double zk = shfl_h(ga, k);
                    const double ua = colA[SM::colL(k)] * inva, ub = colB[SM::colB(k)];
                    ga = fma(-ua, zk, ga);
                    fd = fma(-ub, zk, fd);
                }
                const double* colC = &sm.LC[lrow];
                double gd = fd * invd;
#pragma unroll 4
                for (int k = 0; k < T - 1; ++k) {
                    const double zk = shfl_h(gd, k);
                    const double uc = colC[SM::colL(k)] * invd;
                    gd = fma(-uc, zk, gd);
                }
                fa = ga; fd = gd;                              // = D^-1 L^-1 r
            };

            // [phase: predictor+corrector]
            // ---- predictor (ph = 0) and corrector (ph = 1): one copy of the back solve, the G product and the ratio test ----
            // (unrolled: both passes get their own straight-line code -- no array shuffling through register moves, no
            // `ph` selects: +2 % at T = 20, +3 % at T = 13; it was a rolled loop while the instruction cache was the limiter)
            double dsa[8], dla[8], rc[8];
#pragma unroll
            // no masks in this phase (see mpcb_step.cuh): rows that do not exist on a lane keep s = 1, lam = 0, r_p = 0 by arithmetic
            for (int c = 0; c < 8; ++c) { rc[c] = s[c] * lam[c]; dsa[c] = 0.0; dla[c] = 0.0; }
#pragma unroll
            for (int ph = 0; ph < 2; ++ph) {
                double ha, hd;
                if (ph == 0) { ha = ra * inva; hd = rdl * invd; }              // inva = invd = 0 on the idle lanes
                else {
                    double y[8], ya, yd;
#pragma unroll
                    for (int c = 0; c < 8; ++c) y[c] = rc[c] * sinv[c] - w[c] * rp[c];
                    GT_apply_h<T>(y, dt, l, ya, yd);
                    ha = ya - rda; hd = yd - rdd;
                    fwd_solve(ha, hd);
                }
                double da, dd, g[8];
                back_solve(ha, hd, da, dd);
                G_apply_h<T>(da, dd, dt, l, g);
                double tmax = 0.0;
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    dsa[c] = -rp[c] - g[c];
                    dla[c] = -(rc[c] + lam[c] * dsa[c]) * sinv[c];   // rc = s lam in the predictor: -lam - w ds
                    const double linv = rcp(lam[c]);
                    tmax = dmax_run(dmax_run(tmax, -dsa[c] * sinv[c]), -dla[c] * linv);
                }
                tmax = grp_max_nn(tmax);
                if (ph == 0) {
                    const double al = 1.0 / fmax(1.0, tmax);
                    double sl = 0.0;
#pragma unroll
                    for (int c = 0; c < 8; ++c) sl += (s[c] + al * dsa[c]) * (lam[c] + al * dla[c]);
                    const double mu_aff = grp_sum(sl) / m_cnt;
                    const double r = mu_aff / mu;
                    const double sigma_mu = r * r * r * mu;
                    const double sig_a = mk[0] ? sigma_mu : 0.0, sig_r = mk[4] ? sigma_mu : 0.0;
#pragma unroll
                    for (int c = 0; c < 8; ++c) rc[c] = s[c] * lam[c] + dsa[c] * dla[c] - ((c == 4 || c == 5) ? sig_r : sig_a);
                } else {
                    // Fraction to the boundary: 0.99 while the gap is large, -> 1 as it closes (Mehrotra's adaptive rule).  With a
                    // fixed 0.99 every residual and the gap shrink by exactly 100x per iteration in the end game (linear); with
                    // 1 - mu the end game is quadratic: 2 iterations instead of 4-5 from mu = 1e-4 to 1e-12.  The floor
                    // 0.1 tol_mu / mu keeps one step from driving the gap far below its target (slacks of 1e-20 wreck the
                    // conditioning of H + G'WG while the residuals still need a step).
                    // safeguard: an aggressive step that left the iterate badly centred shows as a gap that fails to halve; from then
                    // on the classic 0.99 (3 solves in 1.3 M stalled without it, profiles/r01g_NOTES.md)
                    if (mu_prev < 0.01 && mu > 0.5 * mu_prev) eta_safe = true;
                    mu_prev = mu;
                    const double eta = eta_safe ? 0.99 : 1.0 - fmin(0.01, fmax(mu, 0.1 * p.tol_mu / mu));
                    const double al = (tmax > eta) ? eta / tmax : 1.0;
                    if (!stop) {                          // a frozen half keeps its iterate (its direction may be NaN)
                        a += al * da; d += al * dd;
                        const double om = 1.0 - al;
#pragma unroll
                        for (int c = 0; c < 8; ++c) { s[c] += al * dsa[c]; lam[c] += al * dla[c]; rp[c] *= om; }
                        rda *= om; rdd *= om; res_d *= om; res_p *= om;
                    }
                }
            }
            if (!stop) { ++it; verified = false; }
            // The carried dual residual is only as good as the backward error of the solves, and a full step makes it exactly
            // zero: from then on the right-hand side no longer corrects the TRUE residual, which drifts while the gap is
            // still closing.  Ordinary problems finish within two iterations of that point (the verification above is their
            // only extra evaluation); a degenerate one that keeps iterating gets a fresh residual every other iteration,
            // while H + G'WG is still well enough conditioned to correct it.
            since = (res_d < p.tol_d + 1e-12 * gscale) ? since + 1 : 0;
            const bool want = !stop && ((it & 15) == 0 || since >= 2);
            if (want) since = 0;
            if (__any_sync(MPCB_FULL, want)) break;
        }
        }
        // [phase: epilogue]
        // OPTIMAL: the configured tolerances, or -- when the loop ended first -- a verified iterate of at least the 1e-8
        // class that the reference's ECOS calls optimal; INACCURATE: ECOS's reduced class (both accepted, mpc.py:196)
        bool ok = !failed && !nearest_fail;
        if (ok) {
            iters = it;
            if (converged) status = MPCB_ST_OPTIMAL;
            else if (feasible0 && best_m < 100.0) {
                status = best_m < 1.0 ? MPCB_ST_OPTIMAL : MPCB_ST_INACCURATE;
                if (act) { a = sm.bc[6][l]; d = sm.bc[7][l]; }
            } else { status = MPCB_ST_FAILED; ok = false; }
        }
        (void)broke;
        if (!ok) { failed = true; solved = false; }      // the lane-per-stage kernel leaves the pass loop here; this half idles on
        else {
            // solution of this pass: mpc.py:197-202
            solved = true;
            oa = a; od = d;
            if (p.lam_ws && act && live) {               // multipliers for the next control step's start (see above)
                double* __restrict__ lw = p.lam_ws + (size_t)b * 8 * T;
#pragma unroll
                for (int c = 0; c < 8; ++c) lw[c * T + l] = lam[c];
            }
        }
        {
            double xs, ys, vs, phs;
            predict(a, d, xs, ys, vs, phs);
            // shift so that l t holds state t; l 0 holds x0
            xs = __shfl_up_sync(MPCB_FULL, xs, 1, MPCB_HW); ys = __shfl_up_sync(MPCB_FULL, ys, 1, MPCB_HW);
            vs = __shfl_up_sync(MPCB_FULL, vs, 1, MPCB_HW); phs = __shfl_up_sync(MPCB_FULL, phs, 1, MPCB_HW);
            if (l == 0) { xs = x0; ys = y0; vs = v0; phs = yaw0; }
            if (ok) { sx = xs; sy = ys; sv = vs; sph = phs; have_ov = true; }
        }
        if (!__any_sync(MPCB_FULL, !failed && !nearest_fail)) break;
    }

    // ---- A7: bookkeeping of MPC.step (main/lib/mpc.py:289-299) -------------------------------------------
    const double od0 = shfl_h(od, 0), oa0 = shfl_h(oa, 0);     // the applied controls (warp-level: before anything diverges)
    if (!live) return 0;                                        // mirrored problem: nothing of it is stored
    if (nearest_fail) {                                         // the reference raises here: nothing of the object changes
        if (l == 0) { p.status[b] = MPCB_ST_NEAREST; if (p.iters) p.iters[b] = 0; if (p.fuse_plant && p.done) p.done[b] = 2; }
        return 0;
    }
    const size_t tb = (size_t)b * 4 * (T + 1);
    if (p.xref && l <= T) {
        p.xref[tb + l] = xr_x;
        p.xref[tb + (T + 1) + l] = xr_y;
        p.xref[tb + 2 * (T + 1) + l] = xr_v;
        p.xref[tb + 3 * (T + 1) + l] = xr_yaw;
    }
    const double qnan = __longlong_as_double(0x7ff8000000000000LL);
    if (solved) {
        if (act) { p.oa[(size_t)b * T + l] = oa; p.od[(size_t)b * T + l] = od; }

        if (p.traj && l <= T) {
            p.traj[tb + l] = sx;
            p.traj[tb + (T + 1) + l] = sy;
            p.traj[tb + 2 * (T + 1) + l] = sv;
            p.traj[tb + 3 * (T + 1) + l] = sph;
        }
    } else if (p.traj && l <= T) {
#pragma unroll
        for (int r = 0; r < 4; ++r) p.traj[tb + r * (T + 1) + l] = qnan;
    }
    if (l == 0) {
        p.target_ind[b] = tind;
        p.status[b] = status;
        if (p.iters) p.iters[b] = iters;
        p.warm[b] = solved ? 1 : 0;
        if (solved) { p.di_ai[2 * b] = od; p.di_ai[2 * b + 1] = oa; }
        else p.di_ai[2 * b + 1] = p.max_decel;                  // mpc.py:296-297, di unchanged
        if (p.xdev) {
            // get_current_xref_deviation (mpc.py:301-308): element-wise product, as written there
            double dev = qnan;
            if (solved && tind < nc) {
                const double ang = cyaw[tind] + 3.141592653589793 / 2;
                const double ex = cos(ang) * (cx[tind] - x0), ey = sin(ang) * (cy[tind] - y0);
                dev = sqrt(ex * ex + ey * ey);
            }
            p.xdev[b] = dev;
        }
    }
    if (p.fuse_plant) {
        // Simulation.step with the applied controls (main/lib/simulation.py:35-47, main/bicycle/main.py:28-41);
        // every l computes the same numbers, lanes 0..7 store one history column each
        const double di = solved ? od0 : p.di_ai[2 * b];
        const double ai = solved ? oa0 : p.max_decel;
        const double dcl = clampd(di, -p.max_steer, p.max_steer);
        double sn, cs;
        sincos(yaw0, &sn, &cs);
        const double xd = __dmul_rn(v0, cs), yd = __dmul_rn(v0, sn);
        const double td = __dmul_rn(__ddiv_rn(v0, L), tan(dcl));
        const double nx = __dadd_rn(x0, __dmul_rn(xd, dt)), ny = __dadd_rn(y0, __dmul_rn(yd, dt));
        const double nyaw = __dadd_rn(yaw0, __dmul_rn(td, dt));
        const double nv = clampd(__dadd_rn(v0, __dmul_rn(ai, dt)), p.min_speed, p.max_speed);
        double dev = qnan;
        if (solved && tind < nc) {
            const double ang = cyaw[tind] + 3.141592653589793 / 2;
            const double ex = cos(ang) * (cx[tind] - x0), ey = sin(ang) * (cy[tind] - y0);
            dev = sqrt(ex * ex + ey * ey);
        }
        double col = nx;
        col = l == 1 ? ny : col;  col = l == 2 ? nyaw : col;  col = l == 3 ? nv : col;
        col = l == 4 ? di : col;  col = l == 5 ? ai : col;    col = l == 6 ? dev : col;
        col = l == 7 ? (double)status : col;
        if (l < 4) p.state_out[4 * b + l] = col;
        if (p.history && l < 8) p.history[8 * (size_t)b + l] = col;
    }
    return iters;
}

// [phase: launch order]
// ---- launch order and kernels: slot s of the launch order is half (s & 1) of warp (s >> 1) -------------------------------
// Neighbouring slots of the longest-first order (lpt_finish, mpcb_step.cuh) took the same number of iterations at the previous
// control step, so the two problems of a warp mostly stop together.
template <int WARPS>
__device__ __forceinline__ int lpt_problem_of_half_slot(const StepParams& p) {
    const int slot = 2 * (blockIdx.x * WARPS + (threadIdx.x >> 5)) + ((threadIdx.x & 31) >> 4);
    if (slot >= p.B) return -1;
    return p.lpt_order ? p.lpt_order[slot] : slot;
}
__device__ __forceinline__ void lpt_finish_half(const StepParams& p, int b, int iters) {
    if (!p.lpt_order) return;
    __shared__ int offs[64];
    const int lane = threadIdx.x & 31;
    int last = 0;
    if ((lane & 15) == 0 && b >= 0) {
        const int bin = 63 - min(max(iters, 0), 63);
        const int r = atomicAdd(&p.lpt_hist[bin], 1);
        p.lpt_rank[b] = (bin << 24) | r;
        __threadfence();
        last = atomicAdd(&p.lpt_hist[64], 1) == p.B - 1;
    }
    if (!__any_sync(MPCB_FULL, last)) return;
    __threadfence();
    const int h0 = __ldcg(&p.lpt_hist[2 * lane]), h1 = __ldcg(&p.lpt_hist[2 * lane + 1]);
    int incl = h0 + h1;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int n = __shfl_up_sync(MPCB_FULL, incl, d);
        if (lane >= d) incl += n;
    }
    offs[2 * lane] = incl - h0 - h1;
    offs[2 * lane + 1] = incl - h1;
    __syncwarp();
    for (int i = lane; i < p.B; i += 32) {
        const int v = __ldcg(&p.lpt_rank[i]);
        p.lpt_order[offs[v >> 24] + (v & 0xffffff)] = i;
    }
    p.lpt_hist[2 * lane] = 0; p.lpt_hist[2 * lane + 1] = 0;
    if (lane == 0) p.lpt_hist[64] = 0;
}

template <int T, int WARPS, int MINB>
__global__ void __launch_bounds__(WARPS * 32, MINB) mpc_step_half_kernel(const StepParams p) {
    const int b = lpt_problem_of_half_slot<WARPS>(p);
    if (__all_sync(MPCB_FULL, b < 0)) return;
    lpt_finish_half(p, b, mpc_step_half_body<T, WARPS, false>(p, b));
}
template <int T, int WARPS, int MINB>
__global__ void __launch_bounds__(WARPS * 32, MINB) mpc_step_half_var_kernel(const StepParams p) {
    const int b = lpt_problem_of_half_slot<WARPS>(p);
    if (__all_sync(MPCB_FULL, b < 0)) return;
    lpt_finish_half(p, b, mpc_step_half_body<T, WARPS, true>(p, b));
}


}  // namespace mpcb
