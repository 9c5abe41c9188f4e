// mpcb_step_long.cuh -- the fused MPC step for LONG horizons (32 <= T <= NT - 1): one CTA per problem,
// thread t <-> horizon stage t, the 2T x 2T system matrix resident in shared memory.
//
// Same reference rows as mpcb_step.cuh (A2..A8: main/lib/mpc.py:86-239,280-326, main/lib/trajectories.py:100-126,
// main/lib/simulation.py:35-47), same condensed model and the same Mehrotra interior point; what changes is the
// execution shape.  BASELINE.json configs[4] sweeps T in {10, 13, 20, 40, 80}: at T = 80 the condensed system has
// n = 160 unknowns, its packed lower triangle is 103 KB -- it cannot live in the registers of one warp, so
//   * every warp-level primitive of the short kernel (scan, neighbour shift, broadcast, reduction) becomes a
//     block-level one (two-level: shuffles inside the warps, one shared-memory pass across them);
//   * M = H + G'WG is assembled every iteration from the per-stage moment tables (H is never stored: an entry is
//     ~10 FMAs from tables that are resident anyway) and factorised in place (right-looking LDL', warps take rows,
//     lanes take columns: conflict-free row-major packed storage);
//   * the gradient is evaluated by the forward/adjoint scans (O(T) per iteration) instead of a Hessian mat-vec.
// This is the throughput-insensitive end of the sweep (n^3 work per iteration): it is built for parity and coverage.
#pragma once
#include "mpcb_common.cuh"
#include "mpcb_step.cuh"

namespace mpcb {

// [phase: block primitives]
// Cross-warp scratch in STATIC shared memory (compile-time addresses: no pointer registers, no address arithmetic), two
// buffers used alternately: primitive k writes buffer k & 1, waits at its one barrier and reads; primitive k + 2 may overwrite
// that buffer because every thread has passed the barrier of primitive k + 1 in between -- one __syncthreads per primitive
// instead of two (the block primitives were 12-18 % of the kernel's samples, all of it barrier wait and address IMADs).
template <int NT>
struct BlockOps {
    static constexpr int NW = NT / 32;
    double (*red)[NW + 1];     // [2][NW + 1]
    int (*redi)[NW + 1];
    int tid, lane, wid, ph;
    __device__ __forceinline__ BlockOps(double (*r)[NW + 1], int (*ri)[NW + 1])
        : red(r), redi(ri), tid(threadIdx.x), lane(threadIdx.x & 31), wid(threadIdx.x >> 5), ph(0) {}
    // inclusive prefix sum over threads 0..tid
    __device__ __forceinline__ double scan_up(double v) {
        double* const r = red[ph ^= 1];
        v = mpcb::scan_up(v, lane);
        if (lane == 31) r[wid] = v;
        __syncthreads();
        double off = 0.0;
#pragma unroll
        for (int w = 0; w < NW; ++w) off += (w < wid) ? r[w] : 0.0;
        return v + off;
    }
    // inclusive suffix sum over threads tid..NT-1
    __device__ __forceinline__ double scan_dn(double v) {
        double* const r = red[ph ^= 1];
        v = mpcb::scan_dn(v, lane);
        if (lane == 0) r[wid] = v;
        __syncthreads();
        double off = 0.0;
#pragma unroll
        for (int w = 0; w < NW; ++w) off += (w > wid) ? r[w] : 0.0;
        return v + off;
    }
    __device__ __forceinline__ double sum(double v) {
        double* const r = red[ph ^= 1];
        v = warp_sum(v);
        if (lane == 0) r[wid] = v;
        __syncthreads();
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < NW; ++w) s += r[w];
        return s;
    }
    __device__ __forceinline__ double max(double v) {
        double* const r = red[ph ^= 1];
        v = warp_max(v);
        if (lane == 0) r[wid] = v;
        __syncthreads();
        double s = r[0];
#pragma unroll
        for (int w = 1; w < NW; ++w) s = fmax(s, r[w]);
        return s;
    }
    // value of thread tid+1 (0 past the end) / tid-1 (0 before the start)
    __device__ __forceinline__ double dn1(double v) {
        double* const r = red[ph ^= 1];
        double n = __shfl_down_sync(MPCB_FULL, v, 1);
        if (lane == 0) r[wid] = v;
        __syncthreads();
        if (lane == 31) n = (wid + 1 < NW) ? r[wid + 1] : 0.0;
        return n;
    }
    __device__ __forceinline__ double up1(double v) {
        double* const r = red[ph ^= 1];
        double n = __shfl_up_sync(MPCB_FULL, v, 1);
        if (lane == 31) r[wid] = v;
        __syncthreads();
        if (lane == 0) n = (wid > 0) ? r[wid - 1] : 0.0;
        return n;
    }
    __device__ __forceinline__ int dn1i(int v) {
        int* const r = redi[ph ^= 1];
        int n = __shfl_down_sync(MPCB_FULL, v, 1);
        if (lane == 0) r[wid] = v;
        __syncthreads();
        if (lane == 31) n = (wid + 1 < NW) ? r[wid + 1] : 0;
        return n;
    }
    // block arg-min on (distance, index), ties to the lower index; every thread gets the winner's index
    __device__ __forceinline__ int argmin(double d, int i) {
        ph ^= 1;
        double* const r = red[ph];
        int* const ri = redi[ph];
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            const double d2 = __shfl_xor_sync(MPCB_FULL, d, o);
            const int i2 = __shfl_xor_sync(MPCB_FULL, i, o);
            if (d2 < d || (d2 == d && i2 < i)) { d = d2; i = i2; }
        }
        if (lane == 0) { r[wid] = d; ri[wid] = i; }
        __syncthreads();
        d = r[0]; i = ri[0];
#pragma unroll
        for (int w = 1; w < NW; ++w) {
            const double d2 = r[w];
            const int i2 = ri[w];
            if (d2 < d || (d2 == d && i2 < i)) { d = d2; i = i2; }
        }
        return i;
    }
};

// main/lib/trajectories.py:100-126, forward=True, block version of nearest_index_in_direction
template <int NT>
__device__ __forceinline__ int nearest_index_block(BlockOps<NT>& bo, const double* __restrict__ cx,
                                                   const double* __restrict__ cy, int start, int nc, double x, double y) {
    const int n = nc - start;
    if (__any_sync(MPCB_FULL, n <= 1)) return start;
    if (__any_sync(MPCB_FULL, n == 2)) return start + 1;
    double bd[3] = {1e300, 1e300, 1e300};
    int bi[3] = {0x7fffffff, 0x7fffffff, 0x7fffffff};
    for (int j = start + bo.tid; j < nc; j += NT) {
        const double dx = __dsub_rn(cx[j], x), dy = __dsub_rn(cy[j], y);
        const double d = __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
        top3_insert(d, j, bd, bi);
    }
    int ind[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        const int i = bo.argmin(bd[0], bi[0]);
        ind[r] = i;
        if (bi[0] == i) {
            bd[0] = bd[1]; bi[0] = bi[1];
            bd[1] = bd[2]; bi[1] = bi[2];
            bd[2] = 1e300; bi[2] = 0x7fffffff;
        }
    }
    if (abs(ind[1] - ind[2]) == 2) return ind[0];
    if (abs(ind[0] - ind[1]) == 1) return max(ind[0], ind[1]);
    return -1;
}

// rank-4 update of N consecutive tiles of one tile row: C_u -= L_I (LD)_u', N tiles in flight (all loads, then the MMAs, then the
// stores: the tiles of a row are consecutive in memory, the compiler cannot prove that a store to one does not feed the next load)
template <int N>
__device__ __forceinline__ void dmma_tiles(const double* __restrict__ wp, double2* cp, const double afrag) {
    double bf[N];
    double2 c[N];
#pragma unroll
    for (int u = 0; u < N; ++u) { bf[u] = wp[32 * u]; c[u] = cp[32 * u]; }
#pragma unroll
    for (int u = 0; u < N; ++u) dmma_m8n8k4(c[u].x, c[u].y, afrag, bf[u]);
#pragma unroll
    for (int u = 0; u < N; ++u) cp[32 * u] = c[u];
}

// shared-memory carve-up of one problem (doubles).  The system matrix M = H + G'WG (n = 2T unknowns, n a multiple of 8) is
// kept as its LOWER 8x8 tiles, tile (I, J <= I) at M + (I (I + 1) / 2 + J) * 64, row-major inside a tile: a tile is the
// accumulator fragment of one mma.m8n8k4 (lane l <-> elements [l/4][2 (l%4)], [l/4][2 (l%4) + 1]: one conflict-free 16-byte
// access per lane), a row chunk of a tile is what one thread of the lane-per-row phases reads.
struct LongSmem {
    // one base pointer and 32-bit offsets (in doubles): the interior-point loop is register-bound, ten generic pointers cost 20
    int T, n, nt;
    double* base;
    int oM, oDinv, oLp, oWp, ovec, odinv, otab;
    __device__ __forceinline__ double* scr() const { return base; }          // jerk variant only: 2 (T + 1)^2 scratch of the pull-back through E
    __device__ __forceinline__ double* M() const { return base + oM; }       // nt (nt + 1) / 2 tiles x 64
    __device__ __forceinline__ double* Dinv() const { return base + oDinv; } // nt x 64: inverses of the unit-lower diagonal tiles of L (zero above the diagonal)
    __device__ __forceinline__ double* Lp() const { return base + oLp; }     // n x 4: -L of the current 4-column panel (rows above the trailing block are zero): A fragments
    __device__ __forceinline__ double* Wp() const { return base + oWp; }     // n x 4:  L D of the current panel (same zero rows): B fragments
    __device__ __forceinline__ double* vec() const { return base + ovec; }   // n      right-hand side / solution
    __device__ __forceinline__ double* dinv() const { return base + odinv; } // n      1 / d_k
    __device__ __forceinline__ double* tab() const { return base + otab; }   // NTAB x (T + 1) per-stage tables
    static constexpr int NTAB = 26;
    __host__ __device__ static size_t unknowns(int T, bool jerk) { return jerk ? (size_t)((2 * T + 1 + 7) & ~7) : 2 * (size_t)T; }
    __host__ __device__ static size_t bytes(int T, int NT, bool jerk = false) {
        const size_t n = unknowns(T, jerk), nt = n / 8;
        (void)NT;
        return (nt * (nt + 1) / 2 * 64 + nt * 64 + 8 * n + 2 * n + (size_t)NTAB * (T + 1) +
                (jerk ? 2 * (size_t)(T + 1) * (T + 1) : 0)) * sizeof(double);
    }
    __device__ __forceinline__ void carve(unsigned char* raw, int T_, int NT, bool jerk = false) {
        T = T_; n = (int)unknowns(T_, jerk); nt = n >> 3;
        base = reinterpret_cast<double*>(raw);
        int q = jerk ? 2 * (T + 1) * (T + 1) : 0;
        oM = q; q += nt * (nt + 1) / 2 * 64;
        oDinv = q; q += nt * 64;
        oLp = q; q += 4 * n;
        oWp = q; q += 4 * n;
        ovec = q; q += n;
        odinv = q; q += n;
        otab = q;
    }
    __device__ __forceinline__ double* t(int i) const { return base + otab + i * (T + 1); }
    __device__ __forceinline__ double* tile(int I, int J) const { return base + oM + ((I * (I + 1) >> 1) + J) * 64; }
};


// table slots
enum { LT_OA = 0, LT_TERM, LT_ZX, LT_ZY, LT_WX, LT_WY, LT_BB, LT_HH, LT_GAX, LT_GAY, LT_GCX, LT_GCY, LT_HA, LT_HC, LT_HB,
       LT_DGA, LT_DGC, LT_WVS, LT_DA, LT_DC, LT_WRP, LT_U0, LT_U1, LT_SPARE, LT_BESTA, LT_BESTD };

// [phase: kernel head/inputs]
// JERK = true: the controller variant main/lib/mpc_jerk.py -- a fifth state z, the integral of the
// acceleration input with a FREE initial value, feeding the speed (`A[2,4] = dt`, :61-86): v_{t+1} = v_t + dt (a_t + z_t).
// With e = a + z ("effective" acceleration), z_k = z_0 + dt sum_{j<k} a_j, the tracking cost and the speed bounds are those of
// the standard model evaluated at e = E [a; z_0]; the kernel pulls gradient, Hessian and speed rows back through E
// (prefix / suffix sums) and carries z_0 as unknown number 2T (n = 2T + 1, padded to a multiple of 8 with identity rows);
// the jerk penalty (z_{t+1} - z_t)^2 = dt^2 a_t^2, t <= T - 2 (:190), is a diagonal input cost.  Derivation checked against the
// goldens recorded from the unmodified module in oracle/condensed_jerk.py.
// One CTA, one problem (b).  Returns the interior-point iterations of the solve (0 if the problem was skipped), block-uniform.
template <int NT, bool JERK>
__device__ __forceinline__ int mpc_step_long_body(const StepParams& p, const int T, const int b) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    LongSmem sm;
    sm.carve(smem_raw, T, NT, JERK);
    __shared__ double s_red[2][NT / 32 + 1];
    __shared__ int s_redi[2][NT / 32 + 1];
    BlockOps<NT> bo(s_red, s_redi);
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    constexpr int NW = NT / 32;
    const bool act = tid < T;
    const double dt = p.dt;
    const int n = sm.n;

    // exits on votes, as in mpcb_step.cuh: block-uniform conditions that the compiler can then also PROVE warp-uniform
    if (__any_sync(MPCB_FULL, p.done && p.done[b] != 0)) return 0;
    const double x0 = p.state[4 * b + 0], y0 = p.state[4 * b + 1], yaw0 = p.state[4 * b + 2], v0 = p.state[4 * b + 3];
    const int route = p.route[b];
    const int nc = p.ncourse[b];
    if (p.fuse_goal) {       // MPC.is_goal (main/lib/mpc.py:310-326)
        const double gd = hypot(x0 - p.goal[2 * b], y0 - p.goal[2 * b + 1]);
        const bool isgoal = gd <= p.goal_dis && abs(p.target_ind[b] - nc) < 5 && fabs(v0) <= p.stop_speed;
        if (__any_sync(MPCB_FULL, isgoal)) { __syncthreads(); if (tid == 0) p.done[b] = 1; return 0; }
    }
    const double L = p.L ? p.L[b] : p.wheelbase;
    const double dl = p.dl[route];
    const double* __restrict__ cx = p.courses + (size_t)route * 3 * p.ncmax;
    const double* __restrict__ cy = cx + p.ncmax;
    const double* __restrict__ cyaw = cy + p.ncmax;
    int tind = p.target_ind[b];
    const int warm = p.warm[b];
    double oa = (warm && act) ? p.oa[(size_t)b * T + tid] : 0.0;     // mpc.py:222-224
    double od = (warm && act) ? p.od[(size_t)b * T + tid] : 0.0;

    double xr_v = 0.0;                                 // xref[2, lane]: speed reference (mpc_with_speed.py:104), else 0
    int status = MPCB_ST_OPTIMAL, iters = 0;
    bool solved = false, have_ov = false;
    int cold_retry = 0;                                // 0: first attempt; else 1 + iterations of the failed attempt from carried multipliers
    double sx = 0.0, sy = 0.0, sv = 0.0, sph = 0.0;    // solution states at index tid (0..T)
    double xr_x = 0.0, xr_y = 0.0, xr_yaw = 0.0;       // xref at state index tid (0..T)

    if (tid == 0) { double* st0 = sm.t(LT_SPARE); st0[0] = x0; st0[1] = y0; st0[2] = yaw0; st0[3] = v0; }   // read by predict (barriers in between)
    for (int pass = 0; pass < p.max_iter; ++pass) {
        // [phase: A2 A3 A4 A5 tables]
        constexpr bool VAR = true;      // the long-horizon kernel always carries the variants (it is not register-bound)
        // cost weights: the handle's, or this problem's own row.  Loaded per pass, not per kernel: ten of the twelve are dead after the
        // A5 tables, and the interior-point loop below is register-bound (values that live across it are spilled in its hot loops)
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
        // ---- A2: reference lookup (main/lib/mpc.py:86-109) ------------------------------------------------
        const int ni = nearest_index_block<NT>(bo, cx, cy, tind, nc, x0, y0);
        if (__any_sync(MPCB_FULL, ni < 0)) { status = MPCB_ST_NEAREST; break; }
        tind = ni;
        double trav = 0.0;
        if (!have_ov) {
            const double pstep = __dmul_rn(fabs(fmax(v0, p.min_ref_speed)), dt);
            double acc = 0.0;
            for (int i = 0; i <= T; ++i) {
                acc = (i == 0) ? pstep : __dadd_rn(acc, pstep);                     // np.cumsum is sequential
                if (i == tid) trav = acc;
            }
        } else {
            if (tid <= T) sm.t(LT_TERM)[tid] = __dmul_rn(fabs(sv), dt);
            __syncthreads();
            double acc = 0.0;
            for (int i = 0; i <= T; ++i) {
                const double ti = sm.t(LT_TERM)[i];
                acc = (i == 0) ? ti : __dadd_rn(acc, ti);
                if (i == tid) trav = acc;
            }
            __syncthreads();
        }
        int ix = (int)rint(__ddiv_rn(trav, dl)) + tind;    // np.rint: half to even
        ix = min(ix, nc - 1);
        if (tid > T) ix = nc - 1;
        ix = max(ix, 0);
        xr_x = cx[ix]; xr_y = cy[ix]; xr_yaw = cyaw[ix];
        xr_v = ix < vcut ? p.ref_speed : 0.0;
        const bool reach = (ix == nc - 1);

        // ---- A3: rollout of the previous (unshifted) solution (main/lib/mpc.py:112-126) ---------------------
        const double dcl = clampd(od, -p.max_steer, p.max_steer);
        const double tnd = act ? tan(dcl) : 0.0;
        if (act) sm.t(LT_OA)[tid] = oa;
        __syncthreads();
        double vb = v0, vcur = v0;
        for (int t = 0; t < T - 1; ++t) {
            vcur = clampd(__dadd_rn(vcur, __dmul_rn(sm.t(LT_OA)[t], dt)), p.min_speed, p.max_speed);
            if (tid == t + 1) vb = vcur;
        }
        const double yterm = __dmul_rn(__dmul_rn(__ddiv_rn(vb, L), tnd), dt);
        if (act) sm.t(LT_TERM)[tid] = yterm;
        __syncthreads();
        double pb = yaw0, pcur = yaw0;
        for (int t = 0; t < T - 1; ++t) {
            pcur = __dadd_rn(pcur, sm.t(LT_TERM)[t]);
            if (tid == t + 1) pb = pcur;
        }
        __syncthreads();

        // ---- A4: linearisation coefficients of stage tid (main/lib/mpc.py:58-79, steer operating point 0) --
        double sn, cs;
        sincos(pb, &sn, &cs);
        const double dtv = dt * vb;
        const double c_ = act ? dt * cs : 0.0, s_ = act ? dt * sn : 0.0;
        const double pp = act ? dtv * sn : 0.0, qq = act ? dtv * cs : 0.0;
        const double bb = act ? dtv / L : 0.0;
        if (!act) pb = 0.0;
        const double cn = bo.dn1(c_), snn = bo.dn1(s_), ppn = bo.dn1(pp), qqn = bo.dn1(qq);

        // ---- A5: stage weights for state tid+1, input weights for stage tid (main/lib/mpc.py:151-184) ------
        const double rx = bo.dn1(xr_x), ry = bo.dn1(xr_y), rpsi = bo.dn1(xr_yaw), rv = bo.dn1(xr_v);
        const bool reach_n = bo.dn1i((int)reach) != 0;
        double m11 = 0.0, m12 = 0.0, m22 = 0.0, wv = 0.0, wp = 0.0;
        if (act) {
            if (!reach_n) {
                double s1, c1, s2, c2;
                sincos(rpsi + 0.5 * 3.141592653589793, &s1, &c1);
                sincos(rpsi, &s2, &c2);
                m11 = w_perp * (c1 * c1) + w_para * (c2 * c2);
                m12 = w_perp * (c1 * s1) + w_para * (c2 * s2);
                m22 = w_perp * (s1 * s1) + w_para * (s2 * s2);
                wv = Qv; wp = Qyaw;
            } else { m11 = Qf0; m22 = Qf1; wv = Qf2; wp = Qf3; }
        }
        const double ra_w = act ? (reach ? p.r_end : R0) : 0.0;
        const double rd_w = act ? (reach ? p.r_end : R1) : 0.0;

        // ---- per-stage moment tables of the condensed Hessian ------------------------------------------------
        const double Zx = bo.scan_up(c_), Zy = bo.scan_up(s_);
        const double Wx = bo.scan_up(-pp), Wy = bo.scan_up(qq);
        {
            const double MZx = m11 * Zx + m12 * Zy, MZy = m12 * Zx + m22 * Zy;
            const double MWx = m11 * Wx + m12 * Wy, MWy = m12 * Wx + m22 * Wy;
            const double S11 = bo.scan_dn(m11), S12 = bo.scan_dn(m12), S22 = bo.scan_dn(m22);
            const double S1zx = bo.scan_dn(MZx), S1zy = bo.scan_dn(MZy);
            const double S1wx = bo.scan_dn(MWx), S1wy = bo.scan_dn(MWy);
            const double S2zz = bo.scan_dn(Zx * MZx + Zy * MZy);
            const double S2ww = bo.scan_dn(Wx * MWx + Wy * MWy);
            const double S2wz = bo.scan_dn(Wx * MZx + Wy * MZy);
            const double Sv = bo.scan_dn(wv), Sp = bo.scan_dn(wp);
            const double gAx = S11 * Zx + S12 * Zy - S1zx, gAy = S12 * Zx + S22 * Zy - S1zy;
            const double gCx = S11 * Wx + S12 * Wy - S1wx, gCy = S12 * Wx + S22 * Wy - S1wy;
            const double hA = S2zz - (Zx * S1zx + Zy * S1zy) + Sv;
            const double hC = S2ww - (Wx * S1wx + Wy * S1wy) + Sp;
            const double hB = S2wz - (Wx * S1zx + Wy * S1zy);
            const double hh = S2wz - (Zx * S1wx + Zy * S1wy);
            const int nb = (tid > 0 ? 1 : 0) + (tid < T - 1 ? 1 : 0);
            if (act) {
                sm.t(LT_ZX)[tid] = Zx; sm.t(LT_ZY)[tid] = Zy; sm.t(LT_WX)[tid] = Wx; sm.t(LT_WY)[tid] = Wy;
                sm.t(LT_BB)[tid] = bb; sm.t(LT_HH)[tid] = hh;
                sm.t(LT_GAX)[tid] = gAx; sm.t(LT_GAY)[tid] = gAy; sm.t(LT_GCX)[tid] = gCx; sm.t(LT_GCY)[tid] = gCy;
                sm.t(LT_HA)[tid] = hA; sm.t(LT_HC)[tid] = hC; sm.t(LT_HB)[tid] = hB;
                sm.t(LT_DGA)[tid] = 2.0 * ra_w + 2.0 * Rd0 * nb; sm.t(LT_DGC)[tid] = 2.0 * rd_w + 2.0 * Rd1 * nb;
            }
            __syncthreads();
        }

        // [phase: predict/grad/G defs]
        // ---- linear prediction and objective gradient by scans (oracle/condensed_model.py states()/grad()) ----
        // jerk variant: effective acceleration e_t = a_t + z_t, z_t = z0 + dt sum_{j<t} a_j, and the adjoint of that map
        // (a gradient with respect to e pulled back to (a, z0))
        auto effective = [&](double a, double z0) {
            if constexpr (!JERK) return a;
            else {
                const double pre = bo.scan_up(a) - a;            // block primitive: every thread takes part
                return act ? a + z0 + dt * pre : 0.0;
            }
        };
        auto pull_back = [&](double ge, double& ga, double& gz) {
            if constexpr (!JERK) { ga = ge; gz = 0.0; }
            else { ga = ge + dt * (bo.scan_dn(ge) - ge); gz = bo.sum(ge); }
        };
        auto predict = [&](double a_in, double d, double z0, double& xs, double& ys, double& vs, double& phs) {
            const double a = effective(a_in, z0);
            const double sa = bo.scan_up(a);
            const double bd_ = bb * d;
            const double sp_ = bo.scan_up(bd_);
            // the initial state from its table row (same values as x0..v0; a broadcast load per use instead of eight registers
            // that would live across the whole interior-point loop)
            const double* const st0 = sm.t(LT_SPARE);
            vs = st0[3] + dt * sa;
            phs = st0[2] + sp_;
            const double vt = st0[3] + dt * (sa - a);
            const double dph = (st0[2] + (sp_ - bd_)) - pb;
            xs = st0[0] + bo.scan_up(c_ * vt - pp * dph);
            ys = st0[1] + bo.scan_up(s_ * vt + qq * dph);
        };
        auto gradient = [&](double a, double d, double z0, double& ga, double& gd, double& gz) {
            double xs, ys, vs, phs;
            predict(a, d, z0, xs, ys, vs, phs);
            const double ex = xs - rx, ey = ys - ry;
            const double lx = 2.0 * (m11 * ex + m12 * ey), ly = 2.0 * (m12 * ex + m22 * ey);
            const double lv = 2.0 * wv * (vs - rv), lp = 2.0 * wp * (phs - rpsi);      // rv = 0: xref[2,:] = 0 (mpc.py:102-105)
            const double px = bo.scan_dn(act ? lx : 0.0), py = bo.scan_dn(act ? ly : 0.0);
            const double pxn = bo.dn1(px), pyn = bo.dn1(py);
            const double pv = bo.scan_dn(act ? lv + cn * pxn + snn * pyn : 0.0);
            const double pph = bo.scan_dn(act ? lp - ppn * pxn + qqn * pyn : 0.0);
            const double ap = bo.up1(a), an = bo.dn1(a);
            const double dp = bo.up1(d), dn = bo.dn1(d);
            const double fa = (tid > 0 ? a - ap : 0.0) - (tid < T - 1 ? an - a : 0.0);
            const double fd = (tid > 0 ? d - dp : 0.0) - (tid < T - 1 ? dn - d : 0.0);
            double gt;
            pull_back(act ? dt * pv : 0.0, gt, gz);
            ga = act ? gt + 2.0 * ra_w * a + 2.0 * Rd0 * fa : 0.0;
            if constexpr (JERK) ga += (tid < T - 1) ? 2.0 * p.jerk_weight * dt * dt * a : 0.0;     // mpc_jerk.py:190
            gd = act ? bb * pph + 2.0 * rd_w * d + 2.0 * Rd1 * fd : 0.0;
        };
        // structured constraint operator, slots as in mpcb_step.cuh (main/lib/mpc.py:184,187-191)
        auto G_apply = [&](double a, double d, double z0, double (&g)[8]) {
            const double dn = bo.dn1(d);
            const double dd = (tid < T - 1) ? dn - d : 0.0;
            const double cs2 = dt * bo.scan_up(effective(a, z0));
            g[0] = a; g[1] = -a; g[2] = d; g[3] = -d; g[4] = dd; g[5] = -dd; g[6] = cs2; g[7] = -cs2;
        };
        auto GT_apply = [&](const double (&y)[8], double& ga, double& gd, double& gz) {
            double gs;
            pull_back(dt * bo.scan_dn(y[6] - y[7]), gs, gz);
            ga = y[0] - y[1] + gs;
            const double r = y[4] - y[5];
            gd = y[2] - y[3] - r + bo.up1(r);
        };

        // [phase: ipm init]
        // ---- A6: interior point on  min J(u)  s.t.  G u <= h -------------------------------------------------
        double hh8[8];
        hh8[0] = p.max_accel; hh8[1] = -p.max_decel; hh8[2] = p.max_steer; hh8[3] = p.max_steer;
        hh8[4] = p.dsteer_dt; hh8[5] = p.dsteer_dt; hh8[6] = p.max_speed - v0; hh8[7] = v0 - p.min_speed;
        bool mk[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) mk[c] = (c == 4 || c == 5) ? (tid < T - 1) : act;
        const double m_cnt = 8.0 * T - 2.0;
        const bool feasible0 = (hh8[6] >= -1e-9) && (hh8[7] >= -1e-9);

        double a = act ? oa : 0.0, d = act ? od : 0.0;
        double z0 = 0.0;                                 // jerk variant: free initial value of the fifth state (every thread holds it)
        double gscale;
        {
            double g0a, g0d, g0z;
            gradient(0.0, 0.0, 0.0, g0a, g0d, g0z);
            gscale = 1.0 + bo.max(fmax(fmax(fabs(g0a), fabs(g0d)), fabs(g0z)));
        }
        double s[8], lam[8];      // centred, gradient-scaled start: see mpcb_step.cuh
        bool carried;
        {
            double g[8];
            G_apply(a, d, z0, g);
            const double mu0 = fmax(10.0, 0.01 * gscale);
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                s[c] = mk[c] ? fmax(hh8[c] - g[c], 1.0) : 1.0;
                lam[c] = mk[c] ? mu0 * rcp(s[c]) : 0.0;
            }
            carried = p.lam_ws && pass == 0 && !cold_retry && p.warm[b];
            if (carried) {                              // multipliers of the previous control step's solve, floored (mpcb_step.cuh)
                const double* __restrict__ lw = p.lam_ws + (size_t)b * 8 * T;
                const double fl = fmax(1.0, 1e-5 * gscale);
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    const double lp = act ? lw[c * T + tid] : 0.0;
                    lam[c] = mk[c] ? fmax(lp, fl) : 0.0;
                }
            }
        }
        double res_d = 1e300, res_p = 1e300, mu = 1e300;
        bool converged = false;
        int it = 0;
        double mu_prev = 1e300;
        bool eta_safe = false;
        // best iterate so far and its merit (see mpcb_step.cuh), kept in two table rows: a_t, d_t at [tid], z0 and the merit at [T]
        // (touched once per iteration; as registers they would be live across the whole loop)
        if (tid == 0) sm.t(LT_BESTD)[T] = 1e300;
        __syncthreads();
        // [phase: residuals + tests]
        while (__any_sync(MPCB_FULL, feasible0)) {
            double rda, rdd, rdz, rp[8];
            {
                double ga, gd, gz, gta, gtd, gtz, g[8];
                gradient(a, d, z0, ga, gd, gz);
                GT_apply(lam, gta, gtd, gtz);
                rda = ga + gta; rdd = gd + gtd; rdz = gz + gtz;
                G_apply(a, d, z0, g);
                double sl = 0.0, rpm = 0.0;
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    rp[c] = mk[c] ? g[c] + s[c] - hh8[c] : 0.0;
                    sl += s[c] * lam[c];
                    rpm = fmax(rpm, fabs(rp[c]));
                }
                mu = bo.sum(sl) / m_cnt;
                res_d = bo.max(fmax(fmax(fabs(rda), fabs(rdd)), fabs(rdz)));
                res_p = bo.max(rpm);
            }
#ifdef MPCB_TRACE
            if (p.trace && tid == 0 && it < 128) {
                double* tr = p.trace + ((size_t)b * 128 + it) * 8;
                tr[0] = it; tr[1] = res_d; tr[2] = res_p; tr[3] = mu; tr[5] = 1.0; tr[7] = gscale;
            }
#endif
            {
                const double m = fmax(fmax(res_d / (1e-8 * gscale), res_p * 1e9), mu * 1e9 / fmax(1.0, 1e-4 * gscale));
                const double bm = sm.t(LT_BESTD)[T];
                __syncthreads();                         // every thread has read the old merit before thread 0 replaces it
                if (m < bm && m < 100.0) {               // block-uniform
                    if (act) { sm.t(LT_BESTA)[tid] = a; sm.t(LT_BESTD)[tid] = d; }
                    if (tid == 0) { sm.t(LT_BESTA)[T] = z0; sm.t(LT_BESTD)[T] = m; }
                }
            }
            if (__any_sync(MPCB_FULL, res_d < p.tol_d + 1e-12 * gscale && res_p < p.tol_p && mu < p.tol_mu)) { converged = true; break; }
            if (__any_sync(MPCB_FULL, it >= p.ipm_max_iter || !(mu == mu) || mu < 1e-15)) break;

            // [phase: W, rhs]
            double sinv[8], w[8];
#pragma unroll
            for (int c = 0; c < 8; ++c) { sinv[c] = rcp(s[c]); w[c] = lam[c] * sinv[c]; }
            {
                const double wv_suf = dt * dt * bo.scan_dn(w[6] + w[7]);
                const double wr = w[4] + w[5];
                const double wr_prev = bo.up1(wr);
                if (act) {
                    sm.t(LT_WVS)[tid] = wv_suf; sm.t(LT_DA)[tid] = w[0] + w[1];
                    sm.t(LT_DC)[tid] = w[2] + w[3] + wr + wr_prev; sm.t(LT_WRP)[tid] = wr_prev;
                }
            }
            // affine right-hand side: -r_d + G'(lam - W r_p)
            {
                double y[8], ya, yd, yz;
#pragma unroll
                for (int c = 0; c < 8; ++c) y[c] = lam[c] - w[c] * rp[c];
                GT_apply(y, ya, yd, yz);
                if (act) { sm.vec()[tid] = ya - rda; sm.vec()[T + tid] = yd - rdd; }
                if constexpr (JERK) { if (tid >= 2 * T && tid < n) sm.vec()[tid] = tid == 2 * T ? yz - rdz : 0.0; }
            }
            __syncthreads();
            // [phase: assemble M]
            // ---- assemble M = H + G'WG tile by tile (unknowns a_0..a_{T-1}, d_0..d_{T-1}; T is a multiple of 4, so a tile
            //      lies inside one of the three blocks): lane l writes elements [l/4][2 (l%4)], [l/4][2 (l%4) + 1] ----------
            if constexpr (JERK) {
                // jerk variant: Hessian in (e, delta) incl. the speed rows' dt^2 L'WL (it acts on e like the tracking cost),
                // pulled back through e = E [a; z0]: P <- E' P E, B <- B E by running sums (T is small: one thread per row /
                // column), then scattered with the input terms into the tiles; unknowns [a | delta | z0 | identity padding]
                const int ld = T + 1;
                double* P = sm.scr();                                // (T + 1) x (T + 1): rows / columns 0..T-1 = e, T = z0
                double* Bq = sm.scr() + ld * ld;                     // T x (T + 1): steer rows, columns as P
                const double kA = 2.0 * dt * dt;
                for (int idx = tid; idx < T * T; idx += NT) {
                    const int i = idx / T, j = idx - i * T, hi = max(i, j), lo = min(i, j);
                    P[i * ld + j] = kA * (sm.t(LT_HA)[hi] + sm.t(LT_ZX)[lo] * sm.t(LT_GAX)[hi] + sm.t(LT_ZY)[lo] * sm.t(LT_GAY)[hi]) + sm.t(LT_WVS)[hi];
                    const double Blo = sm.t(LT_HB)[i] + sm.t(LT_ZX)[j] * sm.t(LT_GCX)[i] + sm.t(LT_ZY)[j] * sm.t(LT_GCY)[i];
                    const double Bup = sm.t(LT_HH)[j] + sm.t(LT_WX)[i] * sm.t(LT_GAX)[j] + sm.t(LT_WY)[i] * sm.t(LT_GAY)[j];
                    Bq[i * ld + j] = 2.0 * sm.t(LT_BB)[i] * dt * ((j <= i) ? Blo : Bup);
                }
                __syncthreads();
                if (tid < 2 * T) {                                 // right-multiply by E: row[k] += dt sum_{t>k} row[t], row[T] = sum
                    double* row = tid < T ? P + tid * ld : Bq + (tid - T) * ld;
                    double tot = 0.0;
                    for (int k = T - 1; k >= 0; --k) { const double v = row[k]; row[k] = v + dt * tot; tot += v; }
                    row[T] = tot;
                }
                __syncthreads();
                if (tid <= T) {                                    // left-multiply P by E': the same down every column
                    double tot = 0.0;
                    for (int k = T - 1; k >= 0; --k) { const double v = P[k * ld + tid]; P[k * ld + tid] = v + dt * tot; tot += v; }
                    P[T * ld + tid] = tot;
                }
                __syncthreads();
                for (int idx = tid; idx < n * n; idx += NT) {
                    const int i = idx / n, j = idx - i * n;
                    if (j > (i | 7)) continue;
                    double x = 0.0;
                    if (j <= i) {
                        if (i < T) {                               // a_i, a_j
                            x = P[i * ld + j];
                            if (j == i) x += sm.t(LT_DGA)[i] + sm.t(LT_DA)[i] + (i < T - 1 ? 2.0 * p.jerk_weight * dt * dt : 0.0);
                            if (j == i - 1) x -= 2.0 * Rd0;
                        } else if (i < 2 * T) {
                            const int r = i - T;
                            if (j < T) x = Bq[r * ld + j];         // delta_r, a_j
                            else {                                 // delta_r, delta_q
                                const int q = j - T;
                                x = 2.0 * sm.t(LT_BB)[r] * sm.t(LT_BB)[q] * (sm.t(LT_HC)[r] + sm.t(LT_WX)[q] * sm.t(LT_GCX)[r] + sm.t(LT_WY)[q] * sm.t(LT_GCY)[r]);
                                if (q == r) x += sm.t(LT_DGC)[r] + sm.t(LT_DC)[r];
                                if (q == r - 1) x -= 2.0 * Rd1 + sm.t(LT_WRP)[r];
                            }
                        } else if (i == 2 * T) {                   // z0 row
                            x = j < T ? P[T * ld + j] : j < 2 * T ? Bq[(j - T) * ld + T] : P[T * ld + T];
                        } else x = (j == i) ? 1.0 : 0.0;           // padding
                    }
                    sm.tile(i >> 3, j >> 3)[(i & 7) * 8 + (j & 7)] = x;
                }
            } else {
                const double kA = 2.0 * dt * dt;
                const int er = lane >> 2, ec = (lane & 3) * 2;
                const int tT = T >> 3;                             // accel tiles per side
                for (int I = wid; I < sm.nt; I += NW) {
                    const int i = 8 * I + er;
                    for (int J = 0; J <= I; ++J) {
                        double v[2];
#pragma unroll
                        for (int e = 0; e < 2; ++e) {
                            const int j = 8 * J + ec + e;
                            double x = 0.0;
                            if (I < tT) {                          // accel row i, accel column j
                                x = kA * (sm.t(LT_HA)[i] + sm.t(LT_ZX)[j] * sm.t(LT_GAX)[i] + sm.t(LT_ZY)[j] * sm.t(LT_GAY)[i]) + sm.t(LT_WVS)[i];
                                if (j == i) x += sm.t(LT_DGA)[i] + sm.t(LT_DA)[i];
                                if (j == i - 1) x -= 2.0 * Rd0;
                            } else {
                                const int r = i - T;
                                const double bbr = sm.t(LT_BB)[r];
                                if (J < tT) {                      // steer row r, accel column j
                                    const double Blo = sm.t(LT_HB)[r] + sm.t(LT_ZX)[j] * sm.t(LT_GCX)[r] + sm.t(LT_ZY)[j] * sm.t(LT_GCY)[r];
                                    const double Bup = sm.t(LT_HH)[j] + sm.t(LT_WX)[r] * sm.t(LT_GAX)[j] + sm.t(LT_WY)[r] * sm.t(LT_GAY)[j];
                                    x = 2.0 * bbr * dt * ((j <= r) ? Blo : Bup);
                                } else {                           // steer row r, steer column q
                                    const int q = j - T;
                                    x = 2.0 * bbr * sm.t(LT_BB)[q] * (sm.t(LT_HC)[r] + sm.t(LT_WX)[q] * sm.t(LT_GCX)[r] + sm.t(LT_WY)[q] * sm.t(LT_GCY)[r]);
                                    if (q == r) x += sm.t(LT_DGC)[r] + sm.t(LT_DC)[r];
                                    if (q == r - 1) x -= 2.0 * Rd1 + sm.t(LT_WRP)[r];
                                }
                            }
                            v[e] = (j <= i) ? x : 0.0;             // strictly upper part of the diagonal tiles: never read
                        }
                        *reinterpret_cast<double2*>(sm.tile(I, J) + er * 8 + ec) = make_double2(v[0], v[1]);
                    }
                }
            }
            // [phase: panel factor]
            // ---- blocked right-looking LDL' in place, panels of 4 columns ------------------------------------------------
            // Per panel: (1) every thread factors the 4x4 pivot block redundantly in registers (broadcast reads, no barrier, no
            // pivot round trip), the thread of row i forward-substitutes its own four entries, stores L in place and drops
            // -L and L D into the fragment buffers; (2) the trailing lower tiles get their rank-4 update on the FP64 tensor
            // cores, M(I,J) -= L(I,:) (L D)(J,:)': one mma.m8n8k4 per tile, accumulator read and written with one 16-byte access
            // per lane, tile rows dealt to the warps in pairs of equal total length.  Two barriers per FOUR pivots.
            bool bad = false;
            for (int k = 0; k < n; k += 4) {
                __syncthreads();
                const int Kt = k >> 3, ko = k & 7;
                const double* Dg = sm.tile(Kt, Kt) + ko * 8 + ko;
                const double m00 = Dg[0], m10 = Dg[8], m11 = Dg[9], m20 = Dg[16], m21 = Dg[17], m22 = Dg[18],
                             m30 = Dg[24], m31 = Dg[25], m32 = Dg[26], m33 = Dg[27];
                const double i0 = rcp(m00);
                const double l10 = m10 * i0, l20 = m20 * i0, l30 = m30 * i0;
                const double d1 = m11 - l10 * m10, w21 = m21 - l20 * m10, w31 = m31 - l30 * m10;
                const double i1 = rcp(d1);
                const double l21 = w21 * i1, l31 = w31 * i1;
                const double d2 = m22 - l20 * m20 - l21 * w21, w32 = m32 - l30 * m20 - l31 * w21;
                const double i2 = rcp(d2);
                const double l32 = w32 * i2;
                const double d3 = m33 - l30 * m30 - l31 * w31 - l32 * w32;
                const double i3 = rcp(d3);
                bad = bad || !(m00 > 0.0 && d1 > 0.0 && d2 > 0.0 && d3 > 0.0 && i0 < 1e300 && i1 < 1e300 && i2 < 1e300 && i3 < 1e300);
                for (int i = tid; i < n; i += NT) {                 // row i of the panel (threads take rows i, i + NT, ...)
                    if (i < k) continue;
                    double2* lp = reinterpret_cast<double2*>(sm.Lp() + 4 * i);
                    double2* wp = reinterpret_cast<double2*>(sm.Wp() + 4 * i);
                    if (i >= k + 4) {
                        double2* row = reinterpret_cast<double2*>(sm.tile(i >> 3, Kt) + (i & 7) * 8 + ko);
                        const double2 r01 = row[0], r23 = row[1];
                        const double w0 = r01.x, l0 = w0 * i0;
                        const double w1 = r01.y - l0 * m10, l1 = w1 * i1;
                        const double w2 = r23.x - l0 * m20 - l1 * w21, l2 = w2 * i2;
                        const double w3 = r23.y - l0 * m30 - l1 * w31 - l2 * w32, l3 = w3 * i3;
                        row[0] = make_double2(l0, l1); row[1] = make_double2(l2, l3);
                        lp[0] = make_double2(-l0, -l1); lp[1] = make_double2(-l2, -l3);
                        wp[0] = make_double2(w0, w1); wp[1] = make_double2(w2, w3);
                    } else {                                       // rows of the pivot block: nothing left to update
                        lp[0] = make_double2(0.0, 0.0); lp[1] = make_double2(0.0, 0.0);
                        wp[0] = make_double2(0.0, 0.0); wp[1] = make_double2(0.0, 0.0);
                        const int a = i - k;
                        sm.dinv()[i] = a == 0 ? i0 : a == 1 ? i1 : a == 2 ? i2 : i3;
                    }
                }
                __syncthreads();
                // [phase: trailing update dmma]
                // the pivot block's own L entries: written now, by the one warp that owns its tile in this phase
                const int Jmin = (k + 4) >> 3;
                const int Rr = sm.nt - Jmin;                       // tile rows (= tile columns) of the trailing block
                auto write_pivot_block = [&]() {
                    double* Dw = sm.tile(Kt, Kt) + ko * 8 + ko;
                    Dw[8] = l10; Dw[16] = l20; Dw[17] = l21; Dw[24] = l30; Dw[25] = l31; Dw[26] = l32;
                };
                if (ko == 4 && tid == 0) write_pivot_block();      // tile (Kt, Kt) is not part of the trailing block
                for (int pr = wid; 2 * pr < Rr; pr += NW) {
#pragma unroll 1
                    for (int half = 0; half < 2; ++half) {
                        const int rel = half == 0 ? pr : Rr - 1 - pr;
                        if (half == 1 && rel == pr) break;
                        const int I = Jmin + rel;
                        const double afrag = sm.Lp()[(8 * I + (lane >> 2)) * 4 + (lane & 3)];
                        // four tiles in flight: all loads, then the MMAs, then the stores (the tiles of a row are consecutive
                        // in memory; the compiler cannot prove that a store to one does not feed the next load)
                        const double* wj = sm.Wp() + (lane >> 2) * 4 + (lane & 3);
                        double2* Crow = reinterpret_cast<double2*>(sm.tile(I, 0) + (lane >> 2) * 8 + (lane & 3) * 2);
                        // tiles Jmin..I of this row, four at a time, then the remainder: fixed offsets inside a group, two pointer
                        // increments per group (the first version clamped and predicated every tile index: 15 instructions per DMMA)
                        const double* wp = wj + 32 * Jmin;
                        double2* cp = Crow + 32 * Jmin;
                        int cnt = I - Jmin + 1;
                        for (; cnt >= 4; cnt -= 4, wp += 128, cp += 128) dmma_tiles<4>(wp, cp, afrag);
                        if (cnt == 3) dmma_tiles<3>(wp, cp, afrag);
                        else if (cnt == 2) dmma_tiles<2>(wp, cp, afrag);
                        else if (cnt == 1) dmma_tiles<1>(wp, cp, afrag);
                        if (ko == 0 && I == Kt) { __syncwarp(); if (lane == 0) write_pivot_block(); }
                    }
                }
            }
            __syncthreads();
#ifdef MPCB_TRACE
            if (p.trace && tid == 0 && it < 128 && bad) p.trace[((size_t)b * 128 + it) * 8 + 6] = 1.0;
#endif
            if (__any_sync(MPCB_FULL, bad)) break;
            // [phase: diag tile inverses]
            // inverses of the unit-lower diagonal tiles (thread 8 J + c: column c of tile J), so that a block step of the
            // triangular solves is an 8-term dot product instead of 8 dependent steps
            for (int i = tid; i < n; i += NT) {
                const int J = i >> 3, c = i & 7;
                const double* Lt = sm.tile(J, J);
                double x[8];
#pragma unroll
                for (int r = 0; r < 8; ++r) {
                    double acc = (r == c) ? 1.0 : 0.0;
#pragma unroll
                    for (int j = 0; j < r; ++j) acc = fma(-Lt[r * 8 + j], x[j], acc);
                    x[r] = acc;
                    sm.Dinv()[J * 64 + r * 8 + c] = acc;             // zero above the diagonal by construction
                }
            }
            __syncthreads();
            // [phase: solves]
            // L D L' x = vec, in place, one block barrier per tile column
            auto solve = [&]() {
                for (int J = 0; J < sm.nt; ++J) {                  // forward: L z = r
                    const int tj = (8 * J) % NT;                   // first of the 8 threads that own tile J's unknowns
                    if (wid == (tj >> 5)) {
                        const bool mine = (tid >> 3) == (tj >> 3);
                        double z = 0.0;
                        if (mine) {
                            const double* Li = sm.Dinv() + J * 64 + (tid & 7) * 8;
                            const double* v = sm.vec() + 8 * J;
#pragma unroll
                            for (int c = 0; c < 8; ++c) z = fma(Li[c], v[c], z);
                        }
                        __syncwarp();
                        if (mine) sm.vec()[8 * J + (tid & 7)] = z;
                    }
                    __syncthreads();
                    for (int i = tid; i < n; i += NT) {
                        if ((i >> 3) <= J) continue;
                        const double* row = sm.tile(i >> 3, J) + (i & 7) * 8;
                        const double* v = sm.vec() + 8 * J;
                        double acc = sm.vec()[i];
#pragma unroll
                        for (int c = 0; c < 8; ++c) acc = fma(-row[c], v[c], acc);
                        sm.vec()[i] = acc;
                    }
                    __syncwarp();
                }
                __syncthreads();
                for (int i = tid; i < n; i += NT) sm.vec()[i] *= sm.dinv()[i];
                __syncthreads();
                for (int J = sm.nt - 1; J >= 0; --J) {             // backward: L' x = y
                    const int tj = (8 * J) % NT;
                    if (wid == (tj >> 5)) {
                        const bool mine = (tid >> 3) == (tj >> 3);
                        double z = 0.0;
                        if (mine) {
                            const double* Lc = sm.Dinv() + J * 64 + (tid & 7);
                            const double* v = sm.vec() + 8 * J;
#pragma unroll
                            for (int c = 0; c < 8; ++c) z = fma(Lc[c * 8], v[c], z);
                        }
                        __syncwarp();
                        if (mine) sm.vec()[8 * J + (tid & 7)] = z;
                    }
                    __syncthreads();
                    for (int i = tid; i < 8 * J; i += NT) {
                        const double* col = sm.tile(J, i >> 3) + (i & 7);
                        const double* v = sm.vec() + 8 * J;
                        double acc = sm.vec()[i];
#pragma unroll
                        for (int c = 0; c < 8; ++c) acc = fma(-col[c * 8], v[c], acc);
                        sm.vec()[i] = acc;
                    }
                    __syncwarp();
                }
                __syncthreads();
            };

            // [phase: predictor]
            // ---- predictor ---------------------------------------------------------------------------------------
            double dsa[8], dla[8];
            double sigma_mu;
            {
                solve();
                const double da = act ? sm.vec()[tid] : 0.0, dd = act ? sm.vec()[T + tid] : 0.0;
                const double dz = JERK ? sm.vec()[2 * T] : 0.0;
                double g[8];
                G_apply(da, dd, dz, g);
                double tmax = 0.0;
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    dsa[c] = mk[c] ? -rp[c] - g[c] : 0.0;
                    dla[c] = mk[c] ? -lam[c] - w[c] * dsa[c] : 0.0;
                    const double linv = mk[c] ? rcp(lam[c]) : 0.0;
                    tmax = dmax_run(dmax_run(tmax, -dsa[c] * sinv[c]), -dla[c] * linv);
                }
                tmax = bo.max(tmax);
                const double al = 1.0 / fmax(1.0, tmax);
                double sl = 0.0;
#pragma unroll
                for (int c = 0; c < 8; ++c) sl += (s[c] + al * dsa[c]) * (lam[c] + al * dla[c]);
                const double mu_aff = bo.sum(sl) / m_cnt;
                const double r = mu_aff / mu;
                sigma_mu = r * r * r * mu;
            }
            // [phase: corrector + step]
            // ---- corrector ---------------------------------------------------------------------------------------
            {
                double rc[8], y[8], ya, yd, yz;
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    rc[c] = mk[c] ? s[c] * lam[c] + dsa[c] * dla[c] - sigma_mu : 0.0;
                    y[c] = rc[c] * sinv[c] - w[c] * rp[c];
                }
                GT_apply(y, ya, yd, yz);
                if (act) { sm.vec()[tid] = ya - rda; sm.vec()[T + tid] = yd - rdd; }
                if constexpr (JERK) { if (tid >= 2 * T && tid < n) sm.vec()[tid] = tid == 2 * T ? yz - rdz : 0.0; }
                __syncthreads();
                solve();
                const double da = act ? sm.vec()[tid] : 0.0, dd = act ? sm.vec()[T + tid] : 0.0;
                const double dz = JERK ? sm.vec()[2 * T] : 0.0;
                double g[8];
                G_apply(da, dd, dz, g);
                double tmax = 0.0;
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    dsa[c] = mk[c] ? -rp[c] - g[c] : 0.0;
                    dla[c] = mk[c] ? -(rc[c] + lam[c] * dsa[c]) * sinv[c] : 0.0;
                    const double linv = mk[c] ? rcp(lam[c]) : 0.0;
                    tmax = dmax_run(dmax_run(tmax, -dsa[c] * sinv[c]), -dla[c] * linv);
                }
                tmax = bo.max(tmax);
                // adaptive fraction to the boundary, see mpcb_step.cuh
                // safeguard: an aggressive step that left the iterate badly centred shows as a gap that fails to halve; from then
                // on the classic 0.99 (3 solves in 1.3 M stalled without it, profiles/r01g_NOTES.md)
                if (mu_prev < 0.01 && mu > 0.5 * mu_prev) eta_safe = true;
                mu_prev = mu;
                const double eta = eta_safe ? 0.99 : 1.0 - fmin(0.01, fmax(mu, 0.1 * p.tol_mu / mu));
                const double al = (tmax > eta) ? eta / tmax : 1.0;
#ifdef MPCB_TRACE
                if (p.trace && tid == 0 && it < 128) p.trace[((size_t)b * 128 + it) * 8 + 4] = al;
#endif
                a += al * da; d += al * dd; z0 += al * dz;
#pragma unroll
                for (int c = 0; c < 8; ++c) { s[c] += al * dsa[c]; lam[c] += al * dla[c]; }
            }
            ++it;
        }
        // [phase: epilogue]
        iters = it + (cold_retry ? cold_retry - 1 : 0);
        if (__any_sync(MPCB_FULL, !converged && !(feasible0 && sm.t(LT_BESTD)[T] < 100.0))) {
            if (__any_sync(MPCB_FULL, carried && feasible0)) {
                // The multipliers carried over from the previous control step save an iteration on almost every solve, but they are
                // the wrong start when the active set flips (a vehicle far off its path whose saturated steering changes sign: every
                // steer-bound multiplier of the old solve is ~1e5 on the wrong side, the iteration unwinds them through W = lam / s
                // ratios of 1e17 and the unpivoted LDL' breaks down before any iterate is acceptable: 2 solves in 720,896 at T = 80,
                // L = 4.0, tests/test_gpu_bench_workloads.py).  Such a solve is repeated from the cold centred start -- the start
                // the CPU port and the reference always use -- by running this pass again (block-uniform: converged / best_m are).
                cold_retry = 1 + it;
                tind = p.target_ind[b];                 // the pass starts from the object's index again (A2 is not idempotent)
                --pass;
                continue;
            }
            status = MPCB_ST_FAILED; solved = false; break;
        }
        if (converged) status = MPCB_ST_OPTIMAL;
        else {
            status = sm.t(LT_BESTD)[T] < 1.0 ? MPCB_ST_OPTIMAL : MPCB_ST_INACCURATE;
            a = act ? sm.t(LT_BESTA)[tid] : 0.0; d = act ? sm.t(LT_BESTD)[tid] : 0.0; z0 = sm.t(LT_BESTA)[T];
        }
        solved = true;
        oa = a; od = d;
        if (p.lam_ws && act) {
            double* __restrict__ lw = p.lam_ws + (size_t)b * 8 * T;
#pragma unroll
            for (int c = 0; c < 8; ++c) lw[c * T + tid] = lam[c];
        }
        {
            double xs, ys, vs, phs;
            predict(a, d, z0, xs, ys, vs, phs);
            sx = bo.up1(xs); sy = bo.up1(ys); sv = bo.up1(vs); sph = bo.up1(phs);    // thread t holds state t
            if (tid == 0) { sx = x0; sy = y0; sv = v0; sph = yaw0; }
        }
        have_ov = true;
    }

    // ---- A7: bookkeeping of MPC.step (main/lib/mpc.py:289-299) -------------------------------------------------
    if (__any_sync(MPCB_FULL, status == MPCB_ST_NEAREST)) {
        if (tid == 0) { p.status[b] = status; if (p.iters) p.iters[b] = 0; if (p.fuse_plant && p.done) p.done[b] = 2; }
        return 0;
    }
    const size_t tb = (size_t)b * 4 * (T + 1);
    if (p.xref && tid <= T) {
        p.xref[tb + tid] = xr_x;
        p.xref[tb + (T + 1) + tid] = xr_y;
        p.xref[tb + 2 * (T + 1) + tid] = xr_v;
        p.xref[tb + 3 * (T + 1) + tid] = xr_yaw;
    }
    const double qnan = __longlong_as_double(0x7ff8000000000000LL);
    if (solved) {
        if (act) { p.oa[(size_t)b * T + tid] = oa; p.od[(size_t)b * T + tid] = od; }
        if (p.traj && tid <= T) {
            p.traj[tb + tid] = sx;
            p.traj[tb + (T + 1) + tid] = sy;
            p.traj[tb + 2 * (T + 1) + tid] = sv;
            p.traj[tb + 3 * (T + 1) + tid] = sph;
        }
    } else if (p.traj && tid <= T) {
        for (int r = 0; r < 4; ++r) p.traj[tb + r * (T + 1) + tid] = qnan;
    }
    // applied controls: first stage of the solution, else the reference's fallback (mpc.py:294-297)
    if (tid == 0) { sm.vec()[0] = od; sm.vec()[1] = oa; }
    __syncthreads();
    const double di = solved ? sm.vec()[0] : p.di_ai[2 * b];
    const double ai = solved ? sm.vec()[1] : p.max_decel;
    double dev = qnan;
    if (solved && tind < nc) {      // get_current_xref_deviation (mpc.py:301-308)
        const double ang = cyaw[tind] + 3.141592653589793 / 2;
        const double ex = cos(ang) * (cx[tind] - x0), ey = sin(ang) * (cy[tind] - y0);
        dev = sqrt(ex * ex + ey * ey);
    }
    __syncthreads();
    if (tid == 0) {
        p.target_ind[b] = tind;
        p.status[b] = status;
        if (p.iters) p.iters[b] = iters;
        p.warm[b] = solved ? 1 : 0;
        if (solved) { p.di_ai[2 * b] = di; p.di_ai[2 * b + 1] = ai; }
        else p.di_ai[2 * b + 1] = p.max_decel;
        if (p.xdev) p.xdev[b] = dev;
    }
    if (p.fuse_plant) {      // Simulation.step (main/lib/simulation.py:35-47, main/bicycle/main.py:28-41)
        const double dcl = clampd(di, -p.max_steer, p.max_steer);
        double sn, cs;
        sincos(yaw0, &sn, &cs);
        const double xd = __dmul_rn(v0, cs), yd = __dmul_rn(v0, sn);
        const double td = __dmul_rn(__ddiv_rn(v0, L), tan(dcl));
        const double nx = __dadd_rn(x0, __dmul_rn(xd, dt)), ny = __dadd_rn(y0, __dmul_rn(yd, dt));
        const double nyaw = __dadd_rn(yaw0, __dmul_rn(td, dt));
        const double nv = clampd(__dadd_rn(v0, __dmul_rn(ai, dt)), p.min_speed, p.max_speed);
        double col = nx;
        col = tid == 1 ? ny : col;  col = tid == 2 ? nyaw : col;  col = tid == 3 ? nv : col;
        col = tid == 4 ? di : col;  col = tid == 5 ? ai : col;    col = tid == 6 ? dev : col;
        col = tid == 7 ? (double)status : col;
        if (tid < 4) p.state_out[4 * b + tid] = col;
        if (p.history && tid < 8) p.history[8 * (size_t)b + tid] = col;
    }
    return iters;
}

// Launch order: longest first, as in the warp-per-problem kernels (lpt_finish, mpcb_step.cuh).  A launch is 7 (T = 40, B = 4096,
// four CTAs per SM) to 28 (T = 80) waves of problems that take 5..35 interior-point iterations; in index order the long ones
// that start late are the tail of the launch.  Block s works on problem lpt_order[s]; warp 0 files the iteration count.
template <int NT, int MINB = 1, bool JERK = false>
__global__ void __launch_bounds__(NT, MINB) mpc_step_long_kernel(const StepParams p, const int T) {
    if (blockIdx.x >= p.B) return;
    const int b = p.lpt_order ? p.lpt_order[blockIdx.x] : (int)blockIdx.x;
    const int iters = mpc_step_long_body<NT, JERK>(p, T, b);
    if (threadIdx.x < 32) lpt_finish(p, b, iters);
}

}  // namespace mpcb
