// mpcb_common.cuh -- parameter block and warp-level helpers shared by the sm_100a kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define MPCB_FULL 0xffffffffu

// Everything a step kernel needs, passed by value (lives in the constant bank of the launch).
struct StepParams {
    int B, R, ncmax;
    const double* courses;   // [R][3][ncmax]  x | y | yaw(smoothed)
    const double* dl;        // [R]
    const double* state;     // [B,4] x y yaw v
    const int* route;        // [B]
    const int* ncourse;      // [B]
    const double* L;         // [B] or nullptr
    int* target_ind;         // [B] io
    int* warm;               // [B] io
    double* oa;              // [B,T] io
    double* od;              // [B,T] io
    double* traj;            // [B,4,T+1] out (may be null)
    double* xref;            // [B,4,T+1] out (may be null)
    double* di_ai;           // [B,2] io
    double* xdev;            // [B] out (may be null)
    int* status;             // [B] out
    int* iters;              // [B] out (may be null)
    // closed-loop fusion (all optional): goal test before the solve, plant step and history after it
    int* done;               // [B] io or null: != 0 -> the problem is skipped (1 = reached goal, 2 = dead)
    const double* goal;      // [B,2] needed when fuse_goal
    double* state_out;       // [B,4] written when fuse_plant (may alias state)
    double* history;         // [B,8] row of this step or null: x y yaw v delta a xref_dev status
    int fuse_goal, fuse_plant;
    double goal_dis, stop_speed;
    // config (main/config/mpc_config.json after main/lib/mpc.py:17-36)
    double w_perp, w_para, R0, R1, Rd0, Rd1, Qv, Qyaw, Qf0, Qf1, Qf2, Qf3;   // Qf* already scaled by T
    double dsteer_dt;        // MAX_DSTEER[rad/s] * dt
    double max_accel, max_decel, max_steer, max_speed, min_speed, r_end, min_ref_speed, dt, wheelbase;
    int max_iter, ipm_max_iter;
    double tol_mu, tol_p, tol_d;
    // controller variants; all off = main/lib/mpc.py
    const double* pw;        // [B,12] or null: per-problem w_perp w_para R0 R1 Rd0 Rd1 Qv Qyaw Qf0..3 (Qf scaled by T) --
                             //   main/lib/mpc_sensitivity.py:150-158 re-reads its JSON on every solve
    const int* vcut;         // [B] or null: main/lib/mpc_with_speed.py:276-282 cutoff_idx (cv[cutoff_idx:] = 0)
    double jerk_weight;      // main/lib/mpc_jerk.py:30,190 jerk_penalty_weight (jerk kernels only)
    double ref_speed;        // cv elsewhere (mpc_with_speed.py:36,280: 25/3.6); xref[2,t] = cv[idx_t] (:104)
    // longest-first launch order for the next control step (resident sessions only; all null = launch slot i solves problem i)
    int* lpt_order;          // [B] io: slot -> problem of THIS launch; rewritten by the last warp to finish, sorted by the
                             //   interior-point iterations each problem just needed (a good predictor of the next step's)
    int* lpt_rank;           // [B] scratch: (bin << 24) | rank inside the bin
    int* lpt_hist;           // [65] scratch: 64 bins (bin 0 = most iterations) + the finished-warp counter; zero between launches
    // interior-point warm start across control steps (resident sessions only): the multipliers of the previous solve
    double* lam_ws;          // [B][8][T] io or null: slot-major rows of the 8 constraint slots per stage
#ifdef MPCB_TRACE
    double* trace;           // development builds only: [B][128][8] per-iteration log (it, res_d, res_p, mu, alpha, flags, ., .)
#endif
};

namespace mpcb {

__device__ __forceinline__ double shfl(double v, int src) { return __shfl_sync(MPCB_FULL, v, src); }
__device__ __forceinline__ double shfl_up1(double v, int lane) {
    double n = __shfl_up_sync(MPCB_FULL, v, 1);
    return lane >= 1 ? n : 0.0;
}
__device__ __forceinline__ double shfl_dn1(double v, int lane) {
    double n = __shfl_down_sync(MPCB_FULL, v, 1);
    return lane < 31 ? n : 0.0;
}
// inclusive prefix sum over lanes 0..lane
__device__ __forceinline__ double scan_up(double v, int lane) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        double n = __shfl_up_sync(MPCB_FULL, v, d);
        if (lane >= d) v += n;
    }
    return v;
}
// inclusive suffix sum over lanes lane..31
__device__ __forceinline__ double scan_dn(double v, int lane) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        double n = __shfl_down_sync(MPCB_FULL, v, d);
        if (lane + d < 32) v += n;
    }
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) v += __shfl_xor_sync(MPCB_FULL, v, d);
    return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) v = fmax(v, __shfl_xor_sync(MPCB_FULL, v, d));
    return v;
}
// reciprocal: hardware seed (23 bits) + one third-order step; operands here are pivots and
// slacks, far from the subnormal range, so the flush-to-zero seed is safe.
__device__ __forceinline__ double rcp(double x) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    // one cubic step r (1 + e + e^2): the seed's relative error e <= 2^-23 leaves e^3 <= 2^-69, in three dependent FMAs
    // instead of the four of two Newton steps (the reciprocal sits on the pivot-to-pivot chain of the LDL')
    const double e = fma(-x, r, 1.0);
    const double t = fma(e, e, e);
    return fma(r, t, r);
}
__device__ __forceinline__ double clampd(double v, double lo, double hi) { return fmax(fmin(v, hi), lo); }
// max(a, b) for a running maximum `a` that is never NaN: a NaN `b` is dropped, as fmax drops it, but in 3 instructions (compare,
// two selects) instead of fmax's 7 (DSETP.MAX, selects, NaN quieting, register moves): the ratio test of the interior point
// takes 32 of them per lane and iteration
__device__ __forceinline__ double dmax_run(double a, double b) { return b > a ? b : a; }
// warp maximum of values none of which is NaN
__device__ __forceinline__ double warp_max_nn(double v) {
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) v = dmax_run(v, __shfl_xor_sync(MPCB_FULL, v, d));
    return v;
}

// Predicated 16-byte shared store (one lane publishes {x, y}); a plain `if (lane == k) ...` with several
// statements compiles to a divergent branch, an unconditional store by all lanes doubles the LSU wavefronts.
__device__ __forceinline__ void sts_pair_if(bool pred, double* ptr, double x, double y) {
    const unsigned addr = static_cast<unsigned>(__cvta_generic_to_shared(ptr));
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %3, 0;\n\t@p st.shared.v2.f64 [%0], {%1, %2};\n\t}"
                 :: "r"(addr), "d"(x), "d"(y), "r"((int)pred) : "memory");
}

// FP64 tensor-core tile: D(8x8) += A(8x4) * B(4x8).  Lane l holds A[l/4][l%4], B[l%4][l/4] and the two
// accumulator entries D[l/4][2*(l%4)], D[l/4][2*(l%4)+1] (PTX mma.sync.aligned.m8n8k4.row.col.f64).
__device__ __forceinline__ void dmma_m8n8k4(double& d0, double& d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

}  // namespace mpcb
