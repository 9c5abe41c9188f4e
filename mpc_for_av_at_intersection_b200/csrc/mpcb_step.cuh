// mpcb_step.cuh -- the fused MPC step: one warp per problem, lane t <-> horizon stage t.
//
// Replaces, per problem, main/lib/mpc.py:211-239 (_iterative_linear_mpc_control):
//   A2  reference lookup      main/lib/mpc.py:86-109, main/lib/trajectories.py:100-126
//   A3  bicycle rollout       main/lib/mpc.py:112-126, main/lib/simulation.py:35-47, main/bicycle/main.py:28-41
//   A4  linearisation         main/lib/mpc.py:58-79 (steer operating point 0)
//   A5  cost assembly         main/lib/mpc.py:151-184, condensed onto the inputs
//   A6  constraints + solve   main/lib/mpc.py:184-208; Mehrotra interior point, LDL' of H + G'WG
//   A7  bookkeeping           main/lib/mpc.py:289-299
//
// Math (derived and checked in oracle/condensed_model.py): inputs u = [a_0..a_{T-1} | d_0..d_{T-1}].
//   v_t   = v_0 + dt * sum_{k<t} a_k                 phi_t = phi_0 + sum_{k<t} b_k d_k,  b_k = dt*vbar_k/L
//   x_t+1 = x_t + c_t v_t - p_t (phi_t - phibar_t)   c = dt cos(phibar), p = dt vbar sin(phibar)
//   y_t+1 = y_t + s_t v_t + q_t (phi_t - phibar_t)   s = dt sin(phibar), q = dt vbar cos(phibar)
// so d pos_t/d a_k = dt (Z_t - Z_{k+1}), d pos_t/d d_k = b_k (W_t - W_{k+1}) with Z, W prefix sums of
// (c,s) and (-p,q); every Hessian entry is a combination of SUFFIX sums of moments of the 2x2 stage
// weights evaluated at max(i,j)+1, and the gradient is one forward + one adjoint scan.  The 8T-2
// inequality rows (accel box, steer box, steer rate, speed bounds as prefix sums) are never stored.
//
// Matrix layout: M = [[A, B'], [B, C]] (A = accel block, C = steer block, B = d/a cross block).
// Lane i keeps row i of A, B and C in registers (3T doubles).  Right-looking LDL': at step k every
// lane publishes its (unscaled) column-k entry to shared memory, which is at the same time the
// broadcast buffer of the rank-1 update and the column-major factor used by the triangular solves.
#pragma once
#include <type_traits>

#include "mpcb_common.cuh"

#define MPCB_ST_OPTIMAL 0
#define MPCB_ST_INACCURATE 1
#define MPCB_ST_FAILED 2
#define MPCB_ST_NEAREST 3

#ifndef MPCB_SOLVE_UNROLL
#define MPCB_SOLVE_UNROLL 4
#endif

namespace mpcb {

constexpr int kSolveUnroll = MPCB_SOLVE_UNROLL;      // unroll factor of the triangular-solve loops

template <int T>
struct WarpSmem {
    static constexpr int N = 2 * T;
    // Every T x T block is stored column-major with a column stride S = 2 * odd and a small per-column shift, chosen
    // so that BOTH access directions are bank-conflict free for 8-byte loads: down a column (consecutive lanes,
    // consecutive words) and along a row (lane i reads column i: i * S + shift(i) covers all 16 8-byte bank pairs
    // over any 16 consecutive i).  Idle lanes alias rows 4..15 (lane & 15) and hit the same words as live lanes.
    //   cost Hessian blocks and the cross factor:   colB(k) = k * SB + ((k >> 3) & 1)
    //   accel / steer factor blocks:                colL(k) = k * SA + (k & 1) + 2 * ((k >> 3) & 1)
    // colL keeps the entry (row k + j, column k) at parity j for every k (aligned 16-byte broadcast loads in a
    // k-rolled loop).  The factor blocks are ZERO on and above the diagonal: the triangular solves then run without
    // a predicate or select (an update with a zero coefficient is the identity).
    __host__ __device__ static constexpr int stride_2odd(int lo) { int s = (lo + 1) & ~1; return (s / 2) % 2 ? s : s + 2; }
    static constexpr int SB = stride_2odd(T + 1);
    static constexpr int SA = stride_2odd(T + 3);
    __host__ __device__ static constexpr int colB(int k) { return k * SB + ((k >> 3) & 1); }
    __host__ __device__ static constexpr int colL(int k) { return k * SA + (k & 1) + 2 * ((k >> 3) & 1); }
    double HA[T * SB + 2];                     // cost Hessian, accel block (full, symmetric)
    double HC[T * SB + 2];                     // steer block (full, symmetric)
    double HB[T * SB + 2];                     // cross block: column j = accel index, row = steer stage
    alignas(16) double LA[T * SA + 36];        // unscaled factor columns of the accel block (strictly lower, rest 0)
    alignas(16) double LB[T * SB + 36];        // cross block: column k (accel pivot k), rows = steer stages
    alignas(16) double LC[T * SA + 36];        // Schur complement (steer block), same layout as LA
    alignas(16) double2 iz[N];                 // {1 / d_k, broadcast slot of the fused forward solve (divided by d_k)}
    double bc[8][T + 1];                       // per-stage broadcast tables used while assembling H
    double tab[6][T];                          // per-stage model coefficients, re-read by the final prediction
};

// ---- k-rolled LDL' building blocks ---------------------------------------------------------------------------
// The instruction cache, not the FP64 pipe, was the first limiter of the fully unrolled factorisation (ncu:
// sm__icc_request_hit_rate 69 %, no_instruction the top stall): a warp keeps its matrix rows in a register window
// that is SHIFTED by the rank-1 update itself (R[j-1] = R[j] - l * c[j]: the FMA writes one slot down), so the
// pivot column is always slot 0 and the loop over pivots k is an ordinary loop.  JMAX = trailing columns kept.
// Critical path of one step: publish column -> barrier -> broadcast loads -> FMAs.  The reciprocal of the NEXT
// pivot is computed by its owner as soon as the first FMA of the update has produced it, i.e. in the shadow of the
// remaining FMAs, and published together with the column (a0 / iv carry it from step to step).
// [phase: ldl accel steps]
template <int T, int JMAX>
__device__ __forceinline__ void ldl_accel_steps(WarpSmem<T>& sm, double (&R)[T], double (&Q)[T], int k0, int k1, int lane,
                                                double wv_suf, double dA, double& ra, double& rdl, bool& bad,
                                                double& inva, double& a0, double& ivm) {
    using SM = WarpSmem<T>;
#pragma unroll 1
    for (int k = k0; k < k1; ++k) {
        const bool mine = lane == k;
        const int col = SM::colL(k);
        const double b0 = Q[0];
        const double a0s = lane > k ? a0 : 0.0;
        // branch-free publish: idle lanes (>= T) write behind the column, into rows of column k + 1 that step k + 1
        // rewrites (or into the pad after the last column); lanes other than k write their own scratch slot
        sm.LA[col + lane] = a0s; sm.LB[SM::colB(k) + lane] = b0;
        sts_pair_if(mine, &sm.iz[k].x, ivm, ra * ivm);
        __syncwarp();
        const double* pc = &sm.LA[col + k];
        const double2 izk = sm.iz[k];
        const double iv = izk.x, zk = izk.y;
        double c[JMAX + 2];
        c[1] = pc[1];
#pragma unroll
        for (int j = 2; j <= JMAX; j += 2) {
            const double2 cc = *reinterpret_cast<const double2*>(pc + j);
            c[j] = cc.x; c[j + 1] = cc.y;
        }
        const double la = a0s * iv, lb = b0 * iv;
        ra = fma(-a0s, zk, ra);
        rdl = fma(-b0, zk, rdl);
        R[0] = R[1] - la * c[1];
        a0 = R[0] + (lane == k + 1 ? dA : 0.0);                    // the diagonal G'WG term joins when the column is published
        ivm = rcp(a0);
        Q[0] = Q[1] - lb * c[1];
#pragma unroll
        for (int j = 2; j <= JMAX; ++j) { R[j - 1] = R[j] - la * c[j]; Q[j - 1] = Q[j] - lb * c[j]; }
    }
}

// [phase: ldl steer steps]
template <int T, int JMAX>
__device__ __forceinline__ void ldl_steer_steps(WarpSmem<T>& sm, double (&C)[T], int k0, int k1, int lane, bool act,
                                                double dC, double wr_prev, double& rdl, bool& bad, double& invd,
                                                double& c0, double& ivm) {
    using SM = WarpSmem<T>;
#pragma unroll 1
    for (int k = k0; k < k1; ++k) {
        const bool mine = lane == k;
        const int col = SM::colL(k);
        const double c0s = lane > k ? c0 : 0.0;
        sm.LC[col + lane] = c0s;
        sts_pair_if(mine, &sm.iz[T + k].x, ivm, rdl * ivm);
        __syncwarp();
        const double* pc = &sm.LC[col + k];
        const double2 izk = sm.iz[T + k];
        const double iv = izk.x, zk = izk.y;
        double c[JMAX + 2];
        c[1] = pc[1];
#pragma unroll
        for (int j = 2; j <= JMAX; j += 2) {
            const double2 cc = *reinterpret_cast<const double2*>(pc + j);
            c[j] = cc.x; c[j + 1] = cc.y;
        }
        const double lc = c0s * iv;
        rdl = fma(-c0s, zk, rdl);
        C[0] = C[1] - lc * c[1];
        c0 = C[0];                                                  // the G'WG terms are already in the window (added to the scratch)
        ivm = rcp(c0);
#pragma unroll
        for (int j = 2; j <= JMAX; ++j) C[j - 1] = C[j] - lc * c[j];
    }
}

// ---- blocked accel + cross phase on the FP64 tensor cores ---------------------------------------------------------
// The accel block A (lower 8x8 tiles) and the cross block B (all tiles) live in mma.m8n8k4 accumulator fragments.
// Per panel of four pivots: (1) the lanes that hold the panel's columns drop them into a [row][4] buffer, (2) in the
// lane-per-row view the 4x4 pivot block is eliminated with shuffles (the predictor's forward solve rides along) and
// the finished columns are published for the triangular solves, (3) the trailing tiles get one rank-4 update each
// from register fragments: A -= L U', B -= L21 U' with L = scaled panel rows, U = unscaled ones.  Compared with the
// lane-per-row rank-1 updates this removes the operand broadcast through the LSU (one wavefront per double) and
// issues 256 FMAs per instruction.
// [phase: accel blocked dmma]
template <int T>
__device__ __forceinline__ void ldl_accel_blocked(WarpSmem<T>& sm, int lane, bool act, double wv_suf, double dA,
                                                  double& ra, double& rdl, bool& bad, double& inva) {
    using SM = WarpSmem<T>;
    constexpr int NTL = (T + 7) / 8, KB = (T + 3) / 4, NTA = NTL * (NTL + 1) / 2, PR = 8 * NTL;
    const int fr = lane >> 2, fk = lane & 3;
    double* const PA = sm.LC;                       // [PR][4] panel rows of A: in entries, out -L (scaled)
    double* const PQ = sm.LC + PR * 4;              // same for the cross block
    double* const PU = sm.LC + 2 * PR * 4;          // unscaled panel rows of A (the B fragments)
    double* const tw = sm.bc[0];                    // per-stage G'WG terms by index
    double* const td = sm.bc[1];
    if (act) { tw[lane] = wv_suf; td[lane] = dA; }
    if (lane < 4 * (PR - T)) {                      // pad rows of the fragment buffers stay zero
        const int o = (T + (lane >> 2)) * 4 + (lane & 3);
        PA[o] = 0.0; PQ[o] = 0.0; PU[o] = 0.0;
    }
    __syncwarp();
    double At[NTA][2], Qt[NTL][NTL][2];
#pragma unroll
    for (int I = 0; I < NTL; ++I)
#pragma unroll
        for (int J = 0; J < NTL; ++J)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int i = 8 * I + fr, j = 8 * J + 2 * fk + e;
                const bool ok = i < T && j < T;
                const int ic = ok ? i : 0, jc = ok ? j : 0;
                Qt[I][J][e] = ok ? sm.HB[SM::colB(jc) + ic] : 0.0;
                if (J <= I) {
                    const int mx = ic > jc ? ic : jc;
                    const double v = sm.HA[SM::colB(jc) + ic] + tw[mx] + (ic == jc ? td[ic] : 0.0);
                    At[I * (I + 1) / 2 + J][e] = ok ? v : 0.0;
                }
            }
    const int row = act ? lane : 0;
#pragma unroll
    for (int p = 0; p < KB; ++p) {
        constexpr int dummy = 0; (void)dummy;
        const int J = p >> 1, h = p & 1;
        const bool holder = (fk >> 1) == h;
        // (1) panel columns -> [row][4]
#pragma unroll
        for (int I = 0; I < NTL; ++I) {
            const int r = 8 * I + fr;
            if (holder && r < T) {
                if (I >= J) *reinterpret_cast<double2*>(&PA[r * 4 + 2 * (fk & 1)]) = make_double2(At[I * (I + 1) / 2 + J][0], At[I * (I + 1) / 2 + J][1]);
                *reinterpret_cast<double2*>(&PQ[r * 4 + 2 * (fk & 1)]) = make_double2(Qt[I][J][0], Qt[I][J][1]);
            }
        }
        __syncwarp();
        // (2) lane-per-row elimination of the 4x4 pivot block
        double a[4], q[4], ln[4], lqn[4], un[4];
        {
            const double2 a01 = *reinterpret_cast<const double2*>(&PA[row * 4]), a23 = *reinterpret_cast<const double2*>(&PA[row * 4 + 2]);
            const double2 q01 = *reinterpret_cast<const double2*>(&PQ[row * 4]), q23 = *reinterpret_cast<const double2*>(&PQ[row * 4 + 2]);
            a[0] = a01.x; a[1] = a01.y; a[2] = a23.x; a[3] = a23.y;
            q[0] = q01.x; q[1] = q01.y; q[2] = q23.x; q[3] = q23.y;
        }
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int k = 4 * p + c;
            if (k < T) {
                const double dk = shfl(a[c], k);
                bad = bad || !(dk > 0.0);
                const double iv = rcp(dk);
                const bool mine = lane == k;
                inva = mine ? iv : inva;
                const double a0s = (act && lane > k) ? a[c] : 0.0;
                const double q0 = act ? q[c] : 0.0;
                const double la = a0s * iv, lq = q0 * iv;
                const double zk = shfl(ra, k) * iv;
                ra = fma(-a0s, zk, ra);
                rdl = fma(-q0, zk, rdl);
#pragma unroll
                for (int c2 = c + 1; c2 < 4; ++c2) {
                    if (4 * p + c2 < T) {
                        const double u = shfl(a[c], 4 * p + c2);
                        a[c2] = fma(-la, u, a[c2]);
                        q[c2] = fma(-lq, u, q[c2]);
                    }
                }
                sm.LA[SM::colL(k) + lane] = a0s; sm.LB[SM::colB(k) + lane] = q0;      // idle lanes: see ldl_accel_steps
                sts_pair_if(mine, &sm.iz[k].x, iv, 0.0);
                ln[c] = -la; lqn[c] = -lq; un[c] = a0s;
            } else { ln[c] = 0.0; lqn[c] = 0.0; un[c] = 0.0; }
        }
        constexpr bool dummy2 = true; (void)dummy2;
        const int J0 = J + h;                          // first column tile that still has live columns
        if (J0 < NTL) {
            if (act) {
                *reinterpret_cast<double2*>(&PA[row * 4]) = make_double2(ln[0], ln[1]); *reinterpret_cast<double2*>(&PA[row * 4 + 2]) = make_double2(ln[2], ln[3]);
                *reinterpret_cast<double2*>(&PQ[row * 4]) = make_double2(lqn[0], lqn[1]); *reinterpret_cast<double2*>(&PQ[row * 4 + 2]) = make_double2(lqn[2], lqn[3]);
                *reinterpret_cast<double2*>(&PU[row * 4]) = make_double2(un[0], un[1]); *reinterpret_cast<double2*>(&PU[row * 4 + 2]) = make_double2(un[2], un[3]);
            }
            __syncwarp();
            // (3) rank-4 updates of the trailing tiles
            double fa[NTL], fq[NTL], fb[NTL];
#pragma unroll
            for (int X = 0; X < NTL; ++X) {
                fa[X] = PA[(8 * X + fr) * 4 + fk]; fq[X] = PQ[(8 * X + fr) * 4 + fk]; fb[X] = PU[(8 * X + fr) * 4 + fk];
            }
#pragma unroll
            for (int Jp = 0; Jp < NTL; ++Jp) {
                if (Jp >= J0) {
#pragma unroll
                    for (int I = 0; I < NTL; ++I) {
                        if (I >= Jp) dmma_m8n8k4(At[I * (I + 1) / 2 + Jp][0], At[I * (I + 1) / 2 + Jp][1], fa[I], fb[Jp]);
                        dmma_m8n8k4(Qt[I][Jp][0], Qt[I][Jp][1], fq[I], fb[Jp]);
                    }
                }
            }
            __syncwarp();
        }
    }
}

// [phase: nearest idx]
// top-3 insertion keeping (d, idx) ascending; ties keep the earlier (lower) index first
__device__ __forceinline__ void top3_insert(double d, int i, double (&bd)[3], int (&bi)[3]) {
    if (d < bd[2]) {
        if (d < bd[1]) {
            bd[2] = bd[1]; bi[2] = bi[1];
            if (d < bd[0]) { bd[1] = bd[0]; bi[1] = bi[0]; bd[0] = d; bi[0] = i; }
            else { bd[1] = d; bi[1] = i; }
        } else { bd[2] = d; bi[2] = i; }
    }
}

// main/lib/trajectories.py:100-126, forward=True.  Returns the new index or -1 ("something wrong").
// Squared distances, no FMA contraction, so that the ordering is the reference's.
__device__ __forceinline__ int nearest_index_in_direction(const double* __restrict__ cx,
                                                          const double* __restrict__ cy, int start, int nc,
                                                          double x, double y, int lane) {
    const int n = nc - start;
    if (__any_sync(MPCB_FULL, n <= 1)) return start;          // votes: see mpc_step_body
    if (__any_sync(MPCB_FULL, n == 2)) return start + 1;
    double bd[3] = {1e300, 1e300, 1e300};
    int bi[3] = {0x7fffffff, 0x7fffffff, 0x7fffffff};
    for (int j = start + lane; j < nc; j += 32) {
        const double dx = __dsub_rn(cx[j], x), dy = __dsub_rn(cy[j], y);
        const double d = __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
        top3_insert(d, j, bd, bi);
    }
    int ind[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        // warp arg-min on (distance, index)
        double d = bd[0];
        int i = bi[0];
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            const double d2 = __shfl_xor_sync(MPCB_FULL, d, o);
            const int i2 = __shfl_xor_sync(MPCB_FULL, i, o);
            if (d2 < d || (d2 == d && i2 < i)) { d = d2; i = i2; }
        }
        ind[r] = i;
        if (bi[0] == i) {   // this lane held the winner: pop it
            bd[0] = bd[1]; bi[0] = bi[1];
            bd[1] = bd[2]; bi[1] = bi[2];
            bd[2] = 1e300; bi[2] = 0x7fffffff;
        }
    }
    if (abs(ind[1] - ind[2]) == 2) return ind[0];
    if (abs(ind[0] - ind[1]) == 1) return max(ind[0], ind[1]);
    return -1;
}

}  // namespace mpcb

namespace mpcb {

// [phase: G ops]
// ---- structured constraint operator ------------------------------------------------------------
// slots per stage t: 0: a<=amax  1: -a<=-amin  2: d<=smax  3: -d<=smax
//                    4: d[t+1]-d[t]<=r  5: -(d[t+1]-d[t])<=r  (t<=T-2)   6: v[t+1]<=vmax  7: -v[t+1]<=-vmin
// main/lib/mpc.py:184,187-191; the v[0] rows are constants (x[:,0]==x0) and are checked separately.
template <int T>
__device__ __forceinline__ void G_apply(double a, double d, double dt, int lane, double (&g)[8]) {
    const double dn = shfl_dn1(d, lane);
    const double dd = (lane < T - 1) ? dn - d : 0.0;
    const double sc = scan_up(a, lane);                              // (the scan is a warp primitive: never inside the select)
    const double cs = lane < T ? dt * sc : 0.0;                      // every row is zero on the idle lanes (a = d = 0 there)
    g[0] = a; g[1] = -a; g[2] = d; g[3] = -d; g[4] = dd; g[5] = -dd; g[6] = cs; g[7] = -cs;
}
template <int T>
__device__ __forceinline__ void GT_apply(const double (&y)[8], double dt, int lane, double& ga, double& gd) {
    ga = y[0] - y[1] + dt * scan_dn(y[6] - y[7], lane);
    const double r = y[4] - y[5];
    gd = y[2] - y[3] - r + shfl_up1(r, lane);
}

// [phase: kernel head/inputs]
// returns the interior-point iterations the problem took (0 if it was skipped)
template <int T, int WARPS, bool VAR>
__device__ __forceinline__ int mpc_step_body(const StepParams& p, const int b) {
    static_assert(T >= 3 && T <= 31, "one warp per problem needs T+1 <= 32");
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    WarpSmem<T>& sm = reinterpret_cast<WarpSmem<T>*>(smem_raw)[wid];
    const bool act = lane < T;
    const double dt = p.dt;

    // ---- problem inputs ------------------------------------------------------------------------
    // Every exit of this function and of its loops is taken on a VOTE (__any_sync): all lanes compute the same condition anyway,
    // but only a vote result is provably warp-uniform to the compiler -- with plain conditions it guarded every group of
    // shuffles below with a divergence check and an out-of-line slow path (24 BRA.DIV, 520 WARPSYNC, 11.9 K instead of
    // ~7 K SASS instructions; the two-per-warp kernel, written with votes from the start, has none).
    if (__any_sync(MPCB_FULL, p.done && p.done[b] != 0)) return 0;
    const double x0 = p.state[4 * b + 0], y0 = p.state[4 * b + 1], yaw0 = p.state[4 * b + 2], v0 = p.state[4 * b + 3];
    const int route = p.route[b];
    const int nc = p.ncourse[b];
    if (p.fuse_goal) {       // MPC.is_goal (main/lib/mpc.py:310-326) as the loop head of mpc_intersection.py:97
        const double gd = hypot(x0 - p.goal[2 * b], y0 - p.goal[2 * b + 1]);
        const bool isgoal = gd <= p.goal_dis && abs(p.target_ind[b] - nc) < 5 && fabs(v0) <= p.stop_speed;
        if (__any_sync(MPCB_FULL, isgoal)) { if (lane == 0) p.done[b] = 1; return 0; }
    }
    const double L = p.L ? p.L[b] : p.wheelbase;
    const double dl = p.dl[route];
    const double* __restrict__ cx = p.courses + (size_t)route * 3 * p.ncmax;
    const double* __restrict__ cy = cx + p.ncmax;
    const double* __restrict__ cyaw = cy + p.ncmax;
    int tind = p.target_ind[b];
    const int warm = p.warm[b];
    double oa = (warm && act) ? p.oa[(size_t)b * T + lane] : 0.0;     // mpc.py:222-224
    double od = (warm && act) ? p.od[(size_t)b * T + lane] : 0.0;

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
    double xr_v = 0.0;                                 // xref[2, lane]: speed reference (mpc_with_speed.py:104), else 0
    int status = MPCB_ST_OPTIMAL, iters = 0;
    bool solved = false, have_ov = false;
    double sx = 0.0, sy = 0.0, sv = 0.0, sph = 0.0;    // solution states at index lane (0..T)
    double xr_x = 0.0, xr_y = 0.0, xr_yaw = 0.0;       // xref at state index lane (0..T)

    for (int pass = 0; pass < p.max_iter; ++pass) {
        // [phase: A2]
        // ---- A2: reference lookup -----------------------------------------------------------------
        const int ni = nearest_index_in_direction(cx, cy, tind, nc, x0, y0, lane);
        if (__any_sync(MPCB_FULL, ni < 0)) { status = MPCB_ST_NEAREST; break; }
        tind = ni;
        double trav = 0.0;
        if (!__any_sync(MPCB_FULL, have_ov)) {
            const double pstep = __dmul_rn(fabs(fmax(v0, p.min_ref_speed)), dt);    // |ov| * dt, mpc.py:96-98
            double acc = 0.0;
#pragma unroll 1
            for (int i = 0; i <= T; ++i) {
                acc = (i == 0) ? pstep : __dadd_rn(acc, pstep);                     // np.cumsum is sequential
                if (i == lane) trav = acc;
            }
        } else {
            const double term = __dmul_rn(fabs(sv), dt);
            double acc = 0.0;
#pragma unroll 1
            for (int i = 0; i <= T; ++i) {
                const double ti = shfl(term, i);
                acc = (i == 0) ? ti : __dadd_rn(acc, ti);
                if (i == lane) trav = acc;
            }
        }
        int idx = (int)rint(__ddiv_rn(trav, dl)) + tind;    // np.rint: half to even
        idx = min(idx, nc - 1);
        if (lane > T) idx = nc - 1;
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
            const double at = shfl(oa, t);
            vcur = clampd(__dadd_rn(vcur, __dmul_rn(at, dt)), p.min_speed, p.max_speed);
            if (lane == t + 1) vb = vcur;
        }
        const double yterm = __dmul_rn(__dmul_rn(__ddiv_rn(vb, L), tnd), dt);
        double pb = yaw0, pcur = yaw0;
#pragma unroll 1
        for (int t = 0; t < T - 1; ++t) {
            pcur = __dadd_rn(pcur, shfl(yterm, t));
            if (lane == t + 1) pb = pcur;
        }

        // [phase: A4/A5]
        // ---- A4: linearisation coefficients of stage `lane` -----------------------------------------
        double sn, cs;
        sincos(pb, &sn, &cs);
        const double dtv = dt * vb;
        const double c_ = act ? dt * cs : 0.0, s_ = act ? dt * sn : 0.0;
        const double pp = act ? dtv * sn : 0.0, qq = act ? dtv * cs : 0.0;
        const double bb = act ? dtv / L : 0.0;
        if (!act) pb = 0.0;
        const double cn = shfl_dn1(c_, lane), snn = shfl_dn1(s_, lane), ppn = shfl_dn1(pp, lane), qqn = shfl_dn1(qq, lane);

        // ---- A5: stage weights for state lane+1, input weights for stage lane -------------------------
        const double rx = shfl_dn1(xr_x, lane), ry = shfl_dn1(xr_y, lane), rpsi = shfl_dn1(xr_yaw, lane);
        double rv = 0.0;
        if constexpr (VAR) rv = shfl_dn1(xr_v, lane);
        const bool reach_n = __shfl_down_sync(MPCB_FULL, (int)reach, 1) != 0;
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
        const double Zx = scan_up(c_, lane), Zy = scan_up(s_, lane);
        const double Wx = scan_up(-pp, lane), Wy = scan_up(qq, lane);
        {
            const double MZx = m11 * Zx + m12 * Zy, MZy = m12 * Zx + m22 * Zy;
            const double MWx = m11 * Wx + m12 * Wy, MWy = m12 * Wx + m22 * Wy;
            const double S11 = scan_dn(m11, lane), S12 = scan_dn(m12, lane), S22 = scan_dn(m22, lane);
            const double S1zx = scan_dn(MZx, lane), S1zy = scan_dn(MZy, lane);
            const double S1wx = scan_dn(MWx, lane), S1wy = scan_dn(MWy, lane);
            const double S2zz = scan_dn(Zx * MZx + Zy * MZy, lane);
            const double S2ww = scan_dn(Wx * MWx + Wy * MWy, lane);
            const double S2wz = scan_dn(Wx * MZx + Wy * MZy, lane);
            const double Sv = scan_dn(wv, lane), Sp = scan_dn(wp, lane);
            const double gAx = S11 * Zx + S12 * Zy - S1zx, gAy = S12 * Zx + S22 * Zy - S1zy;
            const double gCx = S11 * Wx + S12 * Wy - S1wx, gCy = S12 * Wx + S22 * Wy - S1wy;
            const double hA = S2zz - (Zx * S1zx + Zy * S1zy) + Sv;
            const double hC = S2ww - (Wx * S1wx + Wy * S1wy) + Sp;
            const double hB = S2wz - (Wx * S1zx + Wy * S1zy);
            const double hh = S2wz - (Zx * S1wx + Zy * S1wy);
            __syncwarp();
            if (act) {
                sm.bc[0][lane] = Zx; sm.bc[1][lane] = Zy; sm.bc[2][lane] = Wx; sm.bc[3][lane] = Wy;
                sm.bc[4][lane] = bb; sm.bc[5][lane] = hh; sm.bc[6][lane] = gAx; sm.bc[7][lane] = gAy;
            }
            __syncwarp();
            const double kA = 2.0 * dt * dt, kB = 2.0 * bb * dt, kC = 2.0 * bb;
            const int nb = (lane > 0 ? 1 : 0) + (lane < T - 1 ? 1 : 0);
#pragma unroll 1
            for (int j = 0; j < T; ++j) {
                const double zxj = sm.bc[0][j], zyj = sm.bc[1][j], wxj = sm.bc[2][j], wyj = sm.bc[3][j];
                double Aij = kA * (hA + zxj * gAx + zyj * gAy);
                double Cij = kC * sm.bc[4][j] * (hC + wxj * gCx + wyj * gCy);
                const double Blo = hB + zxj * gCx + zyj * gCy;
                const double Bup = sm.bc[5][j] + Wx * sm.bc[6][j] + Wy * sm.bc[7][j];
                const double Bij = kB * ((j <= lane) ? Blo : Bup);
                if (j == lane) { Aij += 2.0 * ra_w + 2.0 * Rd0 * nb; Cij += 2.0 * rd_w + 2.0 * Rd1 * nb; }
                if (j == lane - 1) { Aij -= 2.0 * Rd0; Cij -= 2.0 * Rd1; }
                if (act) {
                    sm.HB[WarpSmem<T>::colB(j) + lane] = Bij;
                    if (j <= lane) {                             // symmetric blocks: both triangles are stored
                        sm.HA[WarpSmem<T>::colB(j) + lane] = Aij; sm.HC[WarpSmem<T>::colB(j) + lane] = Cij;
                        sm.HA[WarpSmem<T>::colB(lane) + j] = Aij; sm.HC[WarpSmem<T>::colB(lane) + j] = Cij;
                    }
                }
            }
            __syncwarp();
        }

        if (act) {
            const double tv[6] = {c_, s_, pp, qq, bb, pb};
#pragma unroll
            for (int i = 0; i < 6; ++i) sm.tab[i][lane] = tv[i];
        }
        __syncwarp();
        auto tld = [&](int i) { return act ? sm.tab[i][lane] : 0.0; };
        // [phase: predict/grad defs]
        // ---- linear prediction and objective gradient by scans ------------------------------------------
        // returns the states at index lane+1 (lanes < T); see oracle/condensed_model.py states()/grad()
        auto predict = [&](double a, double d, double& xs, double& ys, double& vs, double& phs) {
            const double c_ = tld(0), s_ = tld(1), pp = tld(2), qq = tld(3), bb = tld(4), pb = tld(5);
            const double sa = scan_up(a, lane);
            const double bd_ = bb * d;
            const double sp_ = scan_up(bd_, lane);
            vs = v0 + dt * sa;
            phs = yaw0 + sp_;
            const double vt = v0 + dt * (sa - a);
            const double dph = (yaw0 + (sp_ - bd_)) - pb;
            xs = x0 + scan_up(c_ * vt - pp * dph, lane);
            ys = y0 + scan_up(s_ * vt + qq * dph, lane);
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
            const double px = scan_dn(lx, lane), py = scan_dn(ly, lane);
            const double pxn = shfl_dn1(px, lane), pyn = shfl_dn1(py, lane);
            const double pv = scan_dn(lv + cn * pxn + snn * pyn, lane);
            const double pph = scan_dn(lp - ppn * pxn + qqn * pyn, lane);
            const double ap = shfl_up1(a, lane), an = shfl_dn1(a, lane);
            const double dp = shfl_up1(d, lane), dn = shfl_dn1(d, lane);
            const double fa = (lane > 0 ? a - ap : 0.0) - (lane < T - 1 ? an - a : 0.0);
            const double fd = (lane > 0 ? d - dp : 0.0) - (lane < T - 1 ? dn - d : 0.0);
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
        for (int c = 0; c < 8; ++c) mk[c] = (c == 4 || c == 5) ? (lane < T - 1) : act;
        const double m_cnt = 8.0 * T - 2.0;
        // x[2,0] <= MAX_SPEED and >= MIN_SPEED are constant rows (mpc.py:186-188): infeasible if violated
        bool feasible0 = (hh8[6] >= -1e-9) && (hh8[7] >= -1e-9);

        double a = oa, d = od;
        double gscale, g0a, g0d;
        gradient_scan(0.0, 0.0, g0a, g0d);
        gscale = 1.0 + warp_max(fmax(fabs(g0a), fabs(g0d)));
        // Starting point: slacks s = max(h - G u, 1) at the previous solution, multipliers CENTRED on them, lam = mu0 / s,
        // with the gap mu0 = max(10, gscale / 100) tied to the size of the cost gradient.  A vehicle far off its path has
        // gradients of 1e6..1e7 and multipliers of that order at the optimum: from lam = 1 the iteration needs 60-70 tiny
        // steps to grow them (the failures of tests/test_gpu_bench_workloads.py at T = 80), from a scaled start ~20; for
        // ordinary problems (gscale 1e2..1e4) the count is unchanged (numpy twin: 9.24 -> 9.09 per solve on the bench loop).
        double s[8], lam[8];
        {
            double g[8];
            G_apply<T>(a, d, dt, lane, g);
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
                    const double lp = act ? lw[c * T + lane] : 0.0;
                    lam[c] = mk[c] ? fmax(lp, fl) : 0.0;
                }
            }
        }
        // gradient inside the interior-point loop: g0 + H u from the stored cost Hessian (column j of every block is
        // read down the lanes; the transposed cross block along row `lane`)
        const int grow = act ? lane : (T >= 16 ? (lane & 15) : 0);
        auto gradient = [&](double a, double d, double& ga, double& gd) {
            using SMg = WarpSmem<T>;
            double sa = g0a, sd = g0d;
            const double* const hbr = &sm.HB[SMg::colB(grow)];                   // hbr[j] = B(steer j, accel lane)
#pragma unroll 4
            for (int j = 0; j < T; ++j) {
                const double aj = shfl(a, j), dj = shfl(d, j);
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
        // < 100: ECOS's reduced (OPTIMAL_INACCURATE) class, which main/lib/mpc.py:196 accepts too.  The merit is
        // max(r_d / (1e-8 gscale), r_p / 1e-9, mu / (1e-9 max(1, 1e-4 gscale))): dual residual and gap are judged relative to
        // the size of the cost (gscale = 1 + |grad J(0)|_inf), as ECOS judges gap / |objective|: for a vehicle far off its
        // path (gscale 1e5..1e7, multipliers of that order) an ABSOLUTE gap of 1e-7 means slacks of 1e-13 and lam / s ratios
        // of 1e17 -- beyond what an unpivoted LDL' survives (one T = 80 solve in 1.1 M ended FAILED at mu = 2e-7 where the
        // port reported its inaccurate class: scripts/triage_sweep_failures.py).
        // (the merit lives in shared memory too: the kernel sits at the 255-register limit and these values are touched
        // only where the residuals are fresh -- the first iteration, every sixteenth, and the end game)
        double& best_m = sm.bc[5][0];                   // warp-uniform: every lane writes the same value
        best_m = 1e300;
        int since = 0;                                  // iterations since the carried dual residual fell under its tolerance
        // Two nested loops so that the residual evaluation (gradient mat-vec, G and G' scans: ~6 KB of code that runs two to
        // four times per solve) sits OUTSIDE the address range of the hot iteration loop, which is as large as the 32 KB
        // instruction cache: the inner loop leaves ("break") whenever it needs fresh residuals and the outer one re-enters it.
        // [phase: residuals (outer loop)]
        bool stop = !feasible0;
        while (__any_sync(MPCB_FULL, !stop)) {
            {
                double ga, gd, gta, gtd, g[8];
                gradient(a, d, ga, gd);
                GT_apply<T>(lam, dt, lane, gta, gtd);
                rda = ga + gta; rdd = gd + gtd;
                G_apply<T>(a, d, dt, lane, g);
                double rpm = 0.0, sl = 0.0;
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    rp[c] = mk[c] ? g[c] + s[c] - hh8[c] : 0.0;
                    rpm = fmax(rpm, fabs(rp[c]));
                    sl += s[c] * lam[c];
                }
                res_d = warp_max(fmax(fabs(rda), fabs(rdd)));
                res_p = warp_max(rpm);
                verified = true;
                const double m = fmax(fmax(res_d * (1e8 * rcp(gscale)), res_p * 1e9), warp_sum(sl) / m_cnt * (1e9 * rcp(fmax(1.0, 1e-4 * gscale))));
                if (m < best_m && m < 100.0) {
                    best_m = m;
                    if (act) { sm.bc[6][lane] = a; sm.bc[7][lane] = d; }
                }
            }
        // [phase: gap + stopping tests]
        for (;;) {
            {
                double sl = 0.0;
#pragma unroll
                for (int c = 0; c < 8; ++c) sl += s[c] * lam[c];
                mu = warp_sum(sl) / m_cnt;
            }
#ifdef MPCB_TRACE
            if (p.trace && lane == 0 && it < 128) {
                double* tr = p.trace + ((size_t)b * 128 + it) * 8;
                tr[0] = it; tr[1] = res_d; tr[2] = res_p; tr[3] = mu; tr[5] = verified ? 1.0 : 0.0; tr[7] = gscale;
            }
#endif
            if (__any_sync(MPCB_FULL, res_d < p.tol_d + 1e-12 * gscale && res_p < p.tol_p && mu < p.tol_mu)) {
                if (verified) { converged = true; stop = true; }
                break;                                   // not verified: fresh residuals first
            }
            if (__any_sync(MPCB_FULL, broke || it >= p.ipm_max_iter || !(mu == mu) || mu < 1e-15)) {
                if (verified) stop = true;
                break;
            }

            // [phase: W, rhs]
            // ---- M = H + G' W G into registers (rows `lane` of A, B, C) ---------------------------------
            double sinv[8], w[8];
#pragma unroll
            for (int c = 0; c < 8; ++c) { sinv[c] = rcp(s[c]); w[c] = lam[c] * sinv[c]; }
            using SM = WarpSmem<T>;
            double dC, wr_prev, wv_suf, dA;
            {
                wv_suf = dt * dt * scan_dn(w[6] + w[7], lane);                // dt^2 * sum_{t>=lane}
                const double wr = w[4] + w[5];
                wr_prev = shfl_up1(wr, lane);
                dA = w[0] + w[1];
                dC = w[2] + w[3] + wr + wr_prev;
            }
            // affine right-hand side: -r_d + G'(lam - W r_p)
            double ra, rdl;
            {
                double y[8], ya, yd;
#pragma unroll
                for (int c = 0; c < 8; ++c) y[c] = lam[c] - w[c] * rp[c];
                GT_apply<T>(y, dt, lane, ya, yd);
                ra = ya - rda; rdl = yd - rdd;
            }
            // [phase: factor driver+syrk]
            // ---- LDL' of M = H + G'WG, forward solve of the predictor right-hand side fused -------------------
            bool bad = false;
            double inva = 0.0, invd = 0.0;
            {
#ifdef MPCB_BLOCKED_ACCEL
                // measured alternative (profiles/r01b_NOTES.md): 17 % fewer instructions, but the panel elimination puts
                // shuffle -> reciprocal -> shuffle chains and two shared-memory round trips per panel on the critical
                // path; 9 % slower at two resident warps per scheduler
                ldl_accel_blocked<T>(sm, lane, act, wv_suf, dA, ra, rdl, bad, inva);
#else
                double R[T], Q[T];                               // accel-block row and cross-block row of this lane
#pragma unroll
                for (int j = 0; j < T; ++j) {                    // idle lanes read valid memory, their rows are never used
                    R[j] = sm.HA[SM::colB(j) + lane] + wv_suf;   // dt^2 L'WL: entry (i, j <= i) is this lane's suffix sum
                    Q[j] = sm.HB[SM::colB(j) + lane];
                }
                double a0 = R[0] + (lane == 0 ? dA : 0.0);
                double ivm = rcp(a0);
                __syncwarp();
                // three window sizes (measured: 2 segments 8.02 M solves/s, 3: 8.14, 4: 8.12, 6: 8.10 at B = 4096 -- wasted FMAs on
                // dead columns against bytes of hot-loop code)
                constexpr int K1 = T / 3, K2 = (2 * T) / 3;
                ldl_accel_steps<T, T - 1>(sm, R, Q, 0, K1, lane, wv_suf, dA, ra, rdl, bad, inva, a0, ivm);
                ldl_accel_steps<T, T - 1 - K1>(sm, R, Q, K1, K2, lane, wv_suf, dA, ra, rdl, bad, inva, a0, ivm);
                ldl_accel_steps<T, T - 1 - K2>(sm, R, Q, K2, T, lane, wv_suf, dA, ra, rdl, bad, inva, a0, ivm);
#endif
            }
            {
                // steer block: assembled only now (the accel rows are dead), rank-T update C -= L21 D^-1 L21',
                // then its own LDL'
                // The rank-T update runs on the FP64 tensor cores: S = -L21 D^-1 L21' as 8x8 tiles of
                // mma.m8n8k4 (lower tiles only), operands gathered straight from the stored cross factor -- no operand
                // broadcast through the LSU, 256 FMAs per instruction.  The tiles go through a scratch square (the
                // steer factor's storage, dead until the LDL' below) to reach the lane-per-row layout.
                __syncwarp();
                {
                    constexpr int NTL = (T + 7) / 8;                       // 8-row tiles
                    constexpr int KB = (T + 3) / 4;                        // 4-pivot blocks
                    double acc[NTL * (NTL + 1) / 2][2];
#pragma unroll
                    for (int t = 0; t < NTL * (NTL + 1) / 2; ++t) { acc[t][0] = 0.0; acc[t][1] = 0.0; }
                    const int fr = lane >> 2, fk = lane & 3;              // fragment row / k index of this lane
#pragma unroll 1
                    for (int kb = 0; kb < KB; ++kb) {
                        const int k = 4 * kb + fk;
                        const bool kv = (T % 4 == 0) || k < T;
                        const int kc = kv ? k : 0;
                        const double* col = &sm.LB[SM::colB(kc) + fr];
                        const double ivk = sm.iz[kc].x;
                        double fa[NTL], fb[NTL];
#pragma unroll
                        for (int x = 0; x < NTL; ++x) {
                            const double raw = col[8 * x];                 // rows >= T: pad / next column, dead entries only
                            fb[x] = kv ? raw : 0.0;
                            fa[x] = kv ? -raw * ivk : 0.0;
                        }
#pragma unroll
                        for (int I = 0; I < NTL; ++I)
#pragma unroll
                            for (int J = 0; J <= I; ++J) dmma_m8n8k4(acc[I * (I + 1) / 2 + J][0], acc[I * (I + 1) / 2 + J][1], fa[I], fb[J]);
                    }
                    double* S = sm.LC;                                     // [row][SB], rows and columns < T
#pragma unroll
                    for (int I = 0; I < NTL; ++I)
#pragma unroll
                        for (int J = 0; J <= I; ++J) {
                            const int r = 8 * I + fr, c = 8 * J + 2 * fk;
                            if (r < T && c + 1 < T + (T & 1)) *reinterpret_cast<double2*>(&S[r * SM::SB + c]) =
                                make_double2(acc[I * (I + 1) / 2 + J][0], acc[I * (I + 1) / 2 + J][1]);
                        }
                }
                __syncwarp();
                // G'WG of the steer block (diagonal dC, sub-diagonal -wr_prev) goes into this lane's row of the scratch, i.e.
                // into the window before the elimination starts instead of onto the pivot-to-pivot chain of every step
                if (act) {
                    double* Sw = &sm.LC[lane * SM::SB];
                    Sw[lane] += dC;
                    if (lane >= 1) Sw[lane - 1] -= wr_prev;
                }
                double C[T];
                {
                    const double* Sr = &sm.LC[(act ? lane : 0) * SM::SB];
#pragma unroll
                    for (int j = 0; j + 1 < T; j += 2) {                  // entries above the diagonal are never used
                        const double2 sv = *reinterpret_cast<const double2*>(Sr + j);
                        C[j] = sm.HC[SM::colB(j) + lane] + sv.x;
                        C[j + 1] = sm.HC[SM::colB(j + 1) + lane] + sv.y;
                    }
                    if (T & 1) C[T - 1] = sm.HC[SM::colB(T - 1) + lane] + Sr[T - 1];
                }
                __syncwarp();                                              // the scratch is overwritten by the factor below
                double c0 = C[0];
                double ivm = rcp(c0);
                constexpr int K1 = T / 3, K2 = (2 * T) / 3;
                ldl_steer_steps<T, T - 1>(sm, C, 0, K1, lane, act, dC, wr_prev, rdl, bad, invd, c0, ivm);
                ldl_steer_steps<T, T - 1 - K1>(sm, C, K1, K2, lane, act, dC, wr_prev, rdl, bad, invd, c0, ivm);
                ldl_steer_steps<T, T - 1 - K2>(sm, C, K2, T, lane, act, dC, wr_prev, rdl, bad, invd, c0, ivm);
            }
            __syncwarp();
            // every lane's own pivot reciprocals come back from the published {1/d, z} pairs (one load each instead of a
            // select per elimination step); a pivot that was not positive and finite shows in its reciprocal
#ifndef MPCB_BLOCKED_ACCEL
            inva = act ? sm.iz[lane].x : 0.0;
#endif
            invd = act ? sm.iz[T + lane].x : 0.0;
            bad = bad || (act && !(inva > 0.0 && inva < 1e300 && invd > 0.0 && invd < 1e300));
#ifdef MPCB_TRACE
            if (p.trace && lane == 0 && it < 128 && __any_sync(1u, bad)) p.trace[((size_t)b * 128 + it) * 8 + 6] = 1.0;
#endif
            if (__any_sync(MPCB_FULL, bad)) {           // no step from this factorisation: verify the iterate we have, then stop
                broke = true;
                if (verified) stop = true;
                break;
            }
            if (!act) { ra = 0.0; rdl = 0.0; }
            // [phase: solves]
            // Triangular solves with the pivots in rolled loops.  The factor blocks are zero on and above the diagonal,
            // so every step is shuffle -> FMA with no predicate; the running right-hand sides are kept PRE-SCALED by
            // 1/d_i (the coefficient is scaled instead, off the dependency chain), which makes the value a lane
            // broadcasts at its own step the finished unknown.
            const int lrow = act ? lane : (T >= 16 ? (lane & 15) : 0);
            const double* const LAr = &sm.LA[SM::colL(lrow)];        // LAr[k] = (row k, column lane)
            const double* const LCr = &sm.LC[SM::colL(lrow)];
            const double* const LBr = &sm.LB[SM::colB(lrow)];        // LBr[k] = (steer row k, accel column lane)
            auto back_solve = [&](double ha, double hd, double& xa, double& xd) {      // L' x = h, h = D^-1 z
#pragma unroll kSolveUnroll
                for (int k = T - 1; k >= 0; --k) {
                    const double xk = shfl(hd, k);
                    const double uc = LCr[k] * invd, ub = LBr[k] * inva;
                    hd = fma(-uc, xk, hd);
                    ha = fma(-ub, xk, ha);
                }
                xd = hd;
#pragma unroll kSolveUnroll
                for (int k = T - 1; k >= 1; --k) {
                    const double xk = shfl(ha, k);
                    const double ua = LAr[k] * inva;
                    ha = fma(-ua, xk, ha);
                }
                xa = ha;
            };
            auto fwd_solve = [&](double& fa, double& fd) {     // L z = r, factor read from shared memory; returns D^-1-free z
                const double* colA = &sm.LA[lrow];
                const double* colB = &sm.LB[lrow];
                double ga = fa * inva;
#pragma unroll kSolveUnroll
                for (int k = 0; k < T; ++k) {
                    const double zk = shfl(ga, k);
                    const double ua = colA[SM::colL(k)] * inva, ub = colB[SM::colB(k)];
                    ga = fma(-ua, zk, ga);
                    fd = fma(-ub, zk, fd);
                }
                const double* colC = &sm.LC[lrow];
                double gd = fd * invd;
#pragma unroll kSolveUnroll
                for (int k = 0; k < T - 1; ++k) {
                    const double zk = shfl(gd, k);
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
            // No masks in this phase: a row that does not exist on a lane (idle lanes, the last stage's rate rows) holds
            // s = 1, lam = 0, r_p = 0 from the start, G gives 0 there, and then every quantity below stays at its neutral
            // value by arithmetic alone (ds = 0, dlam = -0, products 0); only the centring term needs its two masks.  The
            // reciprocal of lam = 0 is a NaN that the running maximum drops.  (64 double selects per iteration less.)
            for (int c = 0; c < 8; ++c) { rc[c] = s[c] * lam[c]; dsa[c] = 0.0; dla[c] = 0.0; }
#pragma unroll
            for (int ph = 0; ph < 2; ++ph) {
                double ha, hd;
                if (ph == 0) { ha = ra * inva; hd = rdl * invd; }              // inva = invd = 0 on the idle lanes
                else {
                    double y[8], ya, yd;
#pragma unroll
                    for (int c = 0; c < 8; ++c) y[c] = rc[c] * sinv[c] - w[c] * rp[c];
                    GT_apply<T>(y, dt, lane, ya, yd);
                    ha = ya - rda; hd = yd - rdd;
                    fwd_solve(ha, hd);
                }
                double da, dd, g[8];
                back_solve(ha, hd, da, dd);
                G_apply<T>(da, dd, dt, lane, g);
                double tmax = 0.0;
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    dsa[c] = -rp[c] - g[c];
                    dla[c] = -(rc[c] + lam[c] * dsa[c]) * sinv[c];                 // rc = s lam in the predictor: -lam - w ds
                    const double linv = rcp(lam[c]);
                    tmax = dmax_run(dmax_run(tmax, -dsa[c] * sinv[c]), -dla[c] * linv);
                }
                tmax = warp_max_nn(tmax);
                if (ph == 0) {
                    const double al = 1.0 / fmax(1.0, tmax);
                    double sl = 0.0;
#pragma unroll
                    for (int c = 0; c < 8; ++c) sl += (s[c] + al * dsa[c]) * (lam[c] + al * dla[c]);
                    const double mu_aff = warp_sum(sl) / m_cnt;
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
#ifdef MPCB_TRACE
                    if (p.trace && lane == 0 && it < 128) p.trace[((size_t)b * 128 + it) * 8 + 4] = al;
#endif
                    a += al * da; d += al * dd;
                    const double om = 1.0 - al;
#pragma unroll
                    for (int c = 0; c < 8; ++c) { s[c] += al * dsa[c]; lam[c] += al * dla[c]; rp[c] *= om; }
                    rda *= om; rdd *= om; res_d *= om; res_p *= om;
                }
            }
            ++it;
            verified = false;
            // The carried dual residual is only as good as the backward error of the solves, and a full step makes it exactly
            // zero: from then on the right-hand side no longer corrects the TRUE residual, which drifts while the gap is
            // still closing.  Ordinary problems finish within two iterations of that point (the verification above is their
            // only extra evaluation); a degenerate one that keeps iterating gets a fresh residual every other iteration,
            // while H + G'WG is still well enough conditioned to correct it.
            since = (res_d < p.tol_d + 1e-12 * gscale) ? since + 1 : 0;
            if (__any_sync(MPCB_FULL, (it & 15) == 0 || since >= 2)) { since = 0; break; }
        }
        }
        // [phase: epilogue]
        iters = it;
        // OPTIMAL: the configured tolerances, or -- when the loop ended first -- a verified iterate of at least the 1e-8
        // class that the reference's ECOS calls optimal; INACCURATE: ECOS's reduced class (both accepted, mpc.py:196)
        if (__any_sync(MPCB_FULL, !converged && !(feasible0 && best_m < 100.0))) { status = MPCB_ST_FAILED; solved = false; break; }
        if (converged) status = MPCB_ST_OPTIMAL;
        else {
            status = best_m < 1.0 ? MPCB_ST_OPTIMAL : MPCB_ST_INACCURATE;
            if (act) { a = sm.bc[6][lane]; d = sm.bc[7][lane]; }
        }
        (void)broke;
        // solution of this pass: mpc.py:197-202
        solved = true;
        oa = a; od = d;
        if (p.lam_ws && act) {                           // multipliers for the next control step's start (see above)
            double* __restrict__ lw = p.lam_ws + (size_t)b * 8 * T;
#pragma unroll
            for (int c = 0; c < 8; ++c) lw[c * T + lane] = lam[c];
        }
        {
            double xs, ys, vs, phs;
            predict(a, d, xs, ys, vs, phs);
            // shift so that lane t holds state t; lane 0 holds x0
            sx = __shfl_up_sync(MPCB_FULL, xs, 1); sy = __shfl_up_sync(MPCB_FULL, ys, 1);
            sv = __shfl_up_sync(MPCB_FULL, vs, 1); sph = __shfl_up_sync(MPCB_FULL, phs, 1);
            if (lane == 0) { sx = x0; sy = y0; sv = v0; sph = yaw0; }
        }
        have_ov = true;
    }

    // ---- A7: bookkeeping of MPC.step (main/lib/mpc.py:289-299) -------------------------------------------
    if (__any_sync(MPCB_FULL, status == MPCB_ST_NEAREST)) {          // the reference raises here: nothing of the object changes
        if (lane == 0) { p.status[b] = status; if (p.iters) p.iters[b] = 0; if (p.fuse_plant && p.done) p.done[b] = 2; }
        return 0;
    }
    const size_t tb = (size_t)b * 4 * (T + 1);
    if (p.xref && lane <= T) {
        p.xref[tb + lane] = xr_x;
        p.xref[tb + (T + 1) + lane] = xr_y;
        p.xref[tb + 2 * (T + 1) + lane] = xr_v;
        p.xref[tb + 3 * (T + 1) + lane] = xr_yaw;
    }
    const double qnan = __longlong_as_double(0x7ff8000000000000LL);
    if (solved) {
        if (act) { p.oa[(size_t)b * T + lane] = oa; p.od[(size_t)b * T + lane] = od; }

        if (p.traj && lane <= T) {
            p.traj[tb + lane] = sx;
            p.traj[tb + (T + 1) + lane] = sy;
            p.traj[tb + 2 * (T + 1) + lane] = sv;
            p.traj[tb + 3 * (T + 1) + lane] = sph;
        }
    } else if (p.traj && lane <= T) {
#pragma unroll
        for (int r = 0; r < 4; ++r) p.traj[tb + r * (T + 1) + lane] = qnan;
    }
    if (lane == 0) {
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
        // every lane computes the same numbers, lanes 0..7 store one history column each
        const double di = solved ? shfl(od, 0) : p.di_ai[2 * b];
        const double ai = solved ? shfl(oa, 0) : p.max_decel;
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
        col = lane == 1 ? ny : col;  col = lane == 2 ? nyaw : col;  col = lane == 3 ? nv : col;
        col = lane == 4 ? di : col;  col = lane == 5 ? ai : col;    col = lane == 6 ? dev : col;
        col = lane == 7 ? (double)status : col;
        if (lane < 4) p.state_out[4 * b + lane] = col;
        if (p.history && lane < 8) p.history[8 * (size_t)b + lane] = col;
    }
    return iters;
}

// ---- longest-first launch order ---------------------------------------------------------------------------------
// At B = 4096 a launch is 3.5 waves of 148 x 8 resident warps and the problems take 5..20 interior-point iterations: in
// index order the long ones that start late stretch the tail (list-scheduling simulation: makespan 1.58x the ideal, 1.11x
// longest-first).  The iterations a problem needs change little from one control step to the next, so every warp drops its
// count into a 64-bin histogram on its way out, and the LAST warp of the launch turns histogram + per-problem ranks into
// the slot -> problem table of the next launch (a counting sort: one prefix over 64 bins, one coalesced pass over B).
// Nothing numerical depends on the order; no extra launch, no host involvement, capturable.
template <int WARPS>
__device__ __forceinline__ int lpt_problem_of_slot(const StepParams& p, int& slot) {
    slot = blockIdx.x * WARPS + (threadIdx.x >> 5);
    if (slot >= p.B) return -1;
    return p.lpt_order ? p.lpt_order[slot] : slot;
}
__device__ __forceinline__ void lpt_finish(const StepParams& p, int b, int iters) {
    if (!p.lpt_order) return;
    __shared__ int offs[64];
    const int lane = threadIdx.x & 31;
    int last = 0;
    if (lane == 0) {
        const int bin = 63 - min(max(iters, 0), 63);
        const int r = atomicAdd(&p.lpt_hist[bin], 1);
        p.lpt_rank[b] = (bin << 24) | r;
        __threadfence();
        last = atomicAdd(&p.lpt_hist[64], 1) == p.B - 1;
    }
    if (!__any_sync(MPCB_FULL, last)) return;
    __threadfence();
    // exclusive prefix over the 64 bins, two per lane
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

// Register budget: REGS <= 32 is a minimum-resident-CTA hint (__launch_bounds__), REGS > 32 an explicit cap per
// thread (__maxnreg__; ptxas quantises the launch-bounds form much more coarsely than the hardware allocates).
template <int T, int WARPS, int MINB>
__global__ void __launch_bounds__(WARPS * 32, MINB) mpc_step_kernel(const StepParams p) {
    int slot;
    const int b = lpt_problem_of_slot<WARPS>(p, slot);
    if (__any_sync(MPCB_FULL, b < 0)) return;
    lpt_finish(p, b, mpc_step_body<T, WARPS, false>(p, b));
}

template <int T, int WARPS, int REGS>
__global__ void __maxnreg__(REGS) mpc_step_kernel_r(const StepParams p) {
    int slot;
    const int b = lpt_problem_of_slot<WARPS>(p, slot);
    if (__any_sync(MPCB_FULL, b < 0)) return;
    lpt_finish(p, b, mpc_step_body<T, WARPS, false>(p, b));
}

// The controller variants of SURVEY 8f N3 (speed reference of main/lib/mpc_with_speed.py, per-problem cost weights of
// main/lib/mpc_sensitivity.py) as a second instantiation, so that the default kernel's code is untouched.
template <int T, int WARPS, int MINB>
__global__ void __launch_bounds__(WARPS * 32, MINB) mpc_step_var_kernel(const StepParams p) {
    int slot;
    const int b = lpt_problem_of_slot<WARPS>(p, slot);
    if (__any_sync(MPCB_FULL, b < 0)) return;
    lpt_finish(p, b, mpc_step_body<T, WARPS, true>(p, b));
}

}  // namespace mpcb
