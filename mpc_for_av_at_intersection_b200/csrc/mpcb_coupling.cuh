// mpcb_coupling.cuh -- row A9: inter-vehicle coupling, one warp per ego vehicle.
//
// Replaces, per ego and control step, the part of the reference loop body that decides how far the planned
// path may be followed (main/scenarios/mpc_intersection.py:100-139):
//   1. traj_agent_idx <- calc_nearest_index_in_direction on the FULL path      mpc_intersection.py:103-105
//   2. ego "max-accel" prediction = resample_curve(full[idx:], dl_k)            :110-116, lib/trajectories.py:58-86
//   3. constant-input prediction of every obstacle, 35 frames                    lib/moving_obstacles_prediction.py:21-47
//   4. check_collision_moving_cars: circle pairs over +-frame_window offsets     lib/collision_avoidance.py:66-104
//   5. cutoff_idx = max(idx+1, first_index_within_1mm - margin)                  :129-136, collision_avoidance.py:107-119
// Outputs are integers/flags (bit-exact class).  Comparisons use the reference's float64 operation order
// (norm = sqrt(dx*dx + dy*dy), no FMA contraction); cumulative sums are sequential like np.cumsum.
#pragma once
#include "mpcb_common.cuh"

#define MPCB_CP_MAX_OBS 8          // obstacles per ego (config 3: 7 other vehicles)
#define MPCB_CP_FRAMES 35          // len(np.arange(0, 7.0, 0.2))
#define MPCB_CP_MAX_RES 128        // resampled ego frames kept (spacing >= 0.4 m: 60 m path -> < 64)

struct CouplingParams {
    int B, ncmax, vehicles_per_scenario;   // vehicles_per_scenario > 0: obstacles = the other egos of the scenario
    const double* courses;       // [R][3][ncmax]
    const double* pathcc;        // [R][2 circles][2 (x|y)][ncmax] circle centres of every path point (host-computed)
    const double* accel_ramp;    // [ncmax] np.cumsum(zeros + MAX_ACCEL) (host-computed, sequential sums)
    const int* course_len;       // [R]
    const double* dl;            // [R]
    const double* state;         // [B,4] x y yaw v
    const int* route;            // [B]
    const double* obs;           // [B, MPCB_CP_MAX_OBS, 6] x y v yaw a steer (MovingObstacle*.get()), or null
    const int* n_obs;            // [B] or null
    double* script;              // [B, MPCB_CP_MAX_OBS, 8] io or null: scripted cross traffic (MovingObstacleTIntersection,
                                 //   main/lib/moving_obstacles.py:165-231) stepped HERE, once per launch: xc, yc, theta,
                                 //   counter | direction, turning, speed, offset (<= 0: none).  Takes the place of `obs`.
    double* obs_out;             // [B, MPCB_CP_MAX_OBS, 6] or null: what MovingObstacleTIntersection.get() returned this step
    int* done;                   // [B] or null: finished egos keep their cut (and still act as obstacles)
    const double* goal;          // [B,2], target_ind [B]: with fuse_goal the loop-head test MPC.is_goal
    const int* target_ind;       //   (mpc_intersection.py:97, mpc.py:310-326) runs here, against the PREVIOUS cut
    int fuse_goal;
    double goal_dis, stop_speed;
    int* agent_idx;              // [B] io  traj_agent_idx
    int* ncourse;                // [B] io  in: previous cut length (0 = no tmp_trajectory yet), out: new cut length
    int* collided;               // [B] out
    int* coll_idx;               // [B] out  index into full[agent_idx:], -1 if none
    double* coll_xy;             // [B,2] out (NaN if none)
    int* n_res;                  // [B] out  length of the resampled ego prediction
    int* flags;                  // [B] out  bit0: nearest-index inconsistency, bit1: resample overflow
    double circ[2];              // circle centre offsets along the body axis (front, back)
    double radius, wheelbase, dt, max_accel, max_speed;
    int frame_window, n_frames;
};

namespace mpcb {

struct CouplingSmem {
    double cum[32];                                        // chunk of segment lengths, overwritten by their running sum
    int res_idx[MPCB_CP_MAX_RES];                          // path index of every kept ego frame
    double ocx[MPCB_CP_MAX_OBS][2][MPCB_CP_FRAMES];        // obstacle circle centres per frame
    double ocy[MPCB_CP_MAX_OBS][2][MPCB_CP_FRAMES];
};

// exact `sqrt(dx*dx + dy*dy) <= thr` with the square root taken only inside the rounding band
__device__ __forceinline__ bool within(double ax, double ay, double bx, double by, double thr, double thr2) {
    const double dx = __dsub_rn(ax, bx), dy = __dsub_rn(ay, by);
    const double s = __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
    if (s < thr2 * (1.0 - 1e-12)) return true;
    if (s > thr2 * (1.0 + 1e-12)) return false;
    return __dsqrt_rn(s) <= thr;
}

// forward declaration (mpcb_step.cuh)
__device__ __forceinline__ int nearest_index_in_direction(const double* __restrict__ cx, const double* __restrict__ cy,
                                                          int start, int nc, double x, double y, int lane);

template <int WARPS>
__global__ void __launch_bounds__(WARPS * 32) coupling_kernel(const CouplingParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int b = blockIdx.x * WARPS + wid;
    if (b >= p.B) return;
    if (p.done && p.done[b] != 0) return;
    CouplingSmem& sm = reinterpret_cast<CouplingSmem*>(smem_raw)[wid];
    const double x0 = p.state[4 * b], y0 = p.state[4 * b + 1], v0 = p.state[4 * b + 3];
    const int route = p.route[b];
    const int ncf = p.course_len[route];
    if (p.fuse_goal) {
        const int nc_prev = p.ncourse[b] > 0 ? p.ncourse[b] : ncf;      // len(mpc.cx) before this step's cut
        const double gd = hypot(x0 - p.goal[2 * b], y0 - p.goal[2 * b + 1]);
        if (gd <= p.goal_dis && abs(p.target_ind[b] - nc_prev) < 5 && fabs(v0) <= p.stop_speed) {
            if (lane == 0) p.done[b] = 1;
            return;
        }
    }
    const double* __restrict__ cx = p.courses + (size_t)route * 3 * p.ncmax;
    const double* __restrict__ cy = cx + p.ncmax;
    const double* __restrict__ cyaw = cy + p.ncmax;
    const double* __restrict__ pcc = p.pathcc + (size_t)route * 4 * p.ncmax;     // [circle][x|y][ncmax]
    const double qnan = __longlong_as_double(0x7ff8000000000000LL);

    // ---- 1. traj_agent_idx (mpc_intersection.py:103-105) ------------------------------------------------------
    int aidx = p.agent_idx[b];
    const int ncp = p.ncourse[b];
    bool upd = true;
    if (ncp > 0) upd = (cx[aidx] != cx[ncp - 1]) || (cy[aidx] != cy[ncp - 1]) || (cyaw[aidx] != cyaw[ncp - 1]);
    if (upd) {
        const int ni = nearest_index_in_direction(cx, cy, aidx, ncf, x0, y0, lane);
        if (ni < 0) {      // the reference raises (trajectories.py:120): the ego is dead from here on, like MPCB_ST_NEAREST in the step kernel
            if (lane == 0) { if (p.flags) p.flags[b] = 1; if (p.done) p.done[b] = 2; }
            return;
        }
        aidx = ni;
    }
    const int n = ncf - aidx;                                      // len(trajectory_full[traj_agent_idx:]) >= 1
    int flags = 0;

    // ---- 2. ego prediction: resample_curve with per-point spacing (mpc_intersection.py:110-116) ---------------
    int nres = 0;
    {
        const bool ramp = v0 < p.max_speed;
        const double dl_const = __dmul_rn(p.dt, p.max_speed);
        double acc = 0.0;
        long long q_carry = 0;
        for (int base = 0; base < n; base += 32) {
            const int i = base + lane;
            const bool valid = i < n;
            double dist = 0.0;
            if (valid && i > 0) {
                const double dx = __dsub_rn(cx[aidx + i], cx[aidx + i - 1]), dy = __dsub_rn(cy[aidx + i], cy[aidx + i - 1]);
                dist = __dsqrt_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
            }
            __syncwarp();
            sm.cum[lane] = dist;
            __syncwarp();
            // np.cumsum is sequential: every lane runs the same chain of 32 additions and publishes each partial sum
            // (uniform store), then picks its own -- cheaper than a select per step
#pragma unroll 8
            for (int j = 0; j < 32; ++j) {
                acc = __dadd_rn(acc, sm.cum[j]);
                sm.cum[j] = acc;
            }
            __syncwarp();
            const double mycum = sm.cum[lane];
            const double mya = valid ? p.accel_ramp[i] : 0.0;
            const double dl_i = ramp ? __dmul_rn(p.dt, fmin(__dadd_rn(mya, v0), p.max_speed)) : dl_const;
            const long long q = valid ? (long long)floor(__ddiv_rn(mycum, dl_i)) : 0;
            long long qprev = __shfl_up_sync(MPCB_FULL, q, 1);
            if (lane == 0) qprev = q_carry;
            q_carry = __shfl_sync(MPCB_FULL, q, 31);
            const bool keep = valid && (i == 0 || (q - qprev) >= 1 || i == n - 1);
            const unsigned bal = __ballot_sync(MPCB_FULL, keep);
            const int pos = nres + __popc(bal & ((1u << lane) - 1u));
            if (keep) { if (pos < MPCB_CP_MAX_RES) sm.res_idx[pos] = aidx + i; else flags |= 2; }
            nres += __popc(bal);
        }
        flags = __reduce_or_sync(MPCB_FULL, flags);
        if (nres > MPCB_CP_MAX_RES) nres = MPCB_CP_MAX_RES;
    }

    // ---- 3. obstacle predictions (moving_obstacles_prediction.py:21-47), lane o <-> obstacle o ----------------------
    int n_o;
    if (p.vehicles_per_scenario > 0) n_o = p.vehicles_per_scenario - 1;
    else n_o = p.n_obs ? p.n_obs[b] : 0;
    n_o = min(n_o, MPCB_CP_MAX_OBS);
    double ov0 = 0.0, oa0 = 0.0, ox0 = 0.0, oy0 = 0.0;
    if (lane < n_o) {
        double ox, oy, ov, oyaw, oa, ost;
        if (p.vehicles_per_scenario > 0) {          // the other egos of this scenario, constant speed, zero steer
            const int V = p.vehicles_per_scenario, s0 = (b / V) * V, me = b - s0;
            const int other = s0 + (lane < me ? lane : lane + 1);
            ox = p.state[4 * other]; oy = p.state[4 * other + 1]; oyaw = p.state[4 * other + 2]; ov = p.state[4 * other + 3];
            oa = 0.0; ost = 0.0;
        } else if (p.script) {
            // scripted obstacle: get() (moving_obstacles.py:230-232) feeds this step's prediction, then step() (:225-228)
            // moves the model on -- the order of the reference loop (mpc_intersection.py:118-121 before :154-156)
            double* s8 = p.script + ((size_t)b * MPCB_CP_MAX_OBS + lane) * 8;
            const double xc = s8[0], yc = s8[1], th = s8[2], cnt = s8[3], dir = s8[4], turning = s8[5], spd = s8[6], off = s8[7];
            // forward_velocity (:218-223): `offset is None or counter > offset / dt`
            const double fv = (!(off > 0.0) || cnt > __ddiv_rn(off, p.dt)) ? spd : 0.0;
            // steering_angle (:203-216): x_turn = -10 / 12 (:190,195); -np.pi / 2 and 3 * np.pi / 2 in the reference's operation order
            const double kPi = 3.141592653589793;
            double st = 0.0;
            if (turning != 0.0) {
                if (dir > 0.0) { if (xc >= -10.0 && th > __ddiv_rn(-kPi, 2.0)) st = -0.38; }
                else { if (xc <= 12.0 && th < __ddiv_rn(__dmul_rn(3.0, kPi), 2.0)) st = 0.19; }
            }
            ox = xc; oy = yc; ov = fv; oyaw = th; oa = 0.0; ost = st;
            // Bicycle.step(v, delta) (main/bicycle/main.py:28-41), unfused like numpy
            double sn, cs;
            sincos(th, &sn, &cs);
            s8[0] = __dadd_rn(xc, __dmul_rn(__dmul_rn(fv, cs), p.dt));
            s8[1] = __dadd_rn(yc, __dmul_rn(__dmul_rn(fv, sn), p.dt));
            s8[2] = __dadd_rn(th, __dmul_rn(__dmul_rn(__ddiv_rn(fv, p.wheelbase), tan(st)), p.dt));
            s8[3] = cnt + 1.0;
        } else {
            const double* o6 = p.obs + ((size_t)b * MPCB_CP_MAX_OBS + lane) * 6;
            ox = o6[0]; oy = o6[1]; ov = o6[2]; oyaw = o6[3]; oa = o6[4]; ost = o6[5];
        }
        if (p.obs_out) {
            double* w6 = p.obs_out + ((size_t)b * MPCB_CP_MAX_OBS + lane) * 6;
            w6[0] = ox; w6[1] = oy; w6[2] = ov; w6[3] = oyaw; w6[4] = oa; w6[5] = ost;
        }
        // pass A (sequential per obstacle): the yaw and speed sequences.  x, y of frame f use the OLD v / yaw, the
        // yaw update uses the NEW v (moving_obstacles_prediction.py:33-40); yaw_0..yaw_nf go to ocx[o][0][f]
        // (the last one to ocx[o][1][nf-1]) so that pass B can take their sines and cosines in parallel
        const double tn = tan(ost);
        ov0 = ov; oa0 = oa; ox0 = ox; oy0 = oy;
        for (int f = 0; f < p.n_frames; ++f) {
            sm.ocx[lane][0][f] = oyaw;
            ov = __dadd_rn(ov, __dmul_rn(oa, p.dt));
            oyaw = __dadd_rn(oyaw, __dmul_rn(__dmul_rn(__ddiv_rn(ov, p.wheelbase), tn), p.dt));
        }
        sm.ocx[lane][1][p.n_frames - 1] = oyaw;
    }
    __syncwarp();
    // pass B (parallel): sin / cos of all (obstacle, frame) yaws, in place: cos -> ocx, sin -> ocy
    {
        const int nf1 = p.n_frames + 1, total = n_o * nf1;
        for (int idx = lane; idx < total; idx += 32) {
            const int o = idx / nf1, f = idx - o * nf1;
            double* px = f < p.n_frames ? &sm.ocx[o][0][f] : &sm.ocx[o][1][p.n_frames - 1];
            double* py = f < p.n_frames ? &sm.ocy[o][0][f] : &sm.ocy[o][1][p.n_frames - 1];
            double sn, cs;
            sincos(*px, &sn, &cs);
            *px = cs; *py = sn;
        }
    }
    __syncwarp();
    // pass C (sequential per obstacle): positions and circle centres; frame f overwrites slot f after reading it
    double bx0 = 1e300, bx1 = -1e300, by0 = 1e300, by1 = -1e300;       // bounding box of this obstacle's circle centres
    if (lane < n_o) {
        double ox = ox0, oy = oy0, ov = ov0;
        const int nf = p.n_frames;
        for (int f = 0; f < nf; ++f) {
            const double cs = sm.ocx[lane][0][f], sn = sm.ocy[lane][0][f];
            const double cs2 = f + 1 < nf ? sm.ocx[lane][0][f + 1] : sm.ocx[lane][1][nf - 1];
            const double sn2 = f + 1 < nf ? sm.ocy[lane][0][f + 1] : sm.ocy[lane][1][nf - 1];
            ox = __dadd_rn(ox, __dmul_rn(__dmul_rn(ov, cs), p.dt));
            oy = __dadd_rn(oy, __dmul_rn(__dmul_rn(ov, sn), p.dt));
            ov = __dadd_rn(ov, __dmul_rn(oa0, p.dt));
#pragma unroll
            for (int c = 0; c < 2; ++c) {           // trajectories.py:28-34 with y_offset = 0
                const double qx = __dadd_rn(__dmul_rn(cs2, p.circ[c]), ox), qy = __dadd_rn(__dmul_rn(sn2, p.circ[c]), oy);
                sm.ocx[lane][c][f] = qx; sm.ocy[lane][c][f] = qy;
                bx0 = fmin(bx0, qx); bx1 = fmax(bx1, qx); by0 = fmin(by0, qy); by1 = fmax(by1, qy);
            }
        }
    }
    __syncwarp();

    // ---- 4. first colliding pair in the reference's row order: frame, agent circle, obstacle, offset, obstacle circle
    const double thr = 2.0 * p.radius, thr2 = thr * thr;
    const int W = 2 * p.frame_window + 1, nf = p.n_frames;
    bool hit = false;
    double hx = 0.0, hy = 0.0;
    if (n_o > 0) {
        const int n_it = max(nres, nf);
        for (int f = 0; f < n_it && !hit; ++f) {
            const int pidx = sm.res_idx[min(f, nres - 1)];
            const int g = min(f, nf - 1);
            for (int a = 0; a < 2 && !hit; ++a) {
                const double ax = pcc[(2 * a) * p.ncmax + pidx], ay = pcc[(2 * a + 1) * p.ncmax + pidx];
                // obstacles whose circle centres (any frame) can be within reach of this ego circle: a conservative
                // box test per lane/obstacle; the exact test below runs only for those, in ascending obstacle order
                const double mrg = thr * (1.0 + 1e-9);
                unsigned cand = __ballot_sync(MPCB_FULL, lane < n_o && ax >= bx0 - mrg && ax <= bx1 + mrg &&
                                                             ay >= by0 - mrg && ay <= by1 + mrg);
                while (cand && !hit) {
                    const int o = __ffs(cand) - 1;
                    cand &= cand - 1;
                    for (int base = 0; base < 2 * W; base += 32) {
                        const int idx = base + lane;
                        bool h = false;
                        double px = 0.0, py = 0.0;
                        if (idx < 2 * W) {
                            const int k = idx >> 1, c = idx & 1;
                            const int src = min(max(g - (k - p.frame_window), 0), nf - 1);
                            px = sm.ocx[o][c][src]; py = sm.ocy[o][c][src];
                            h = within(ax, ay, px, py, thr, thr2);
                        }
                        const unsigned bal = __ballot_sync(MPCB_FULL, h);
                        if (bal) {
                            const int src_lane = __ffs(bal) - 1;
                            hx = shfl(px, src_lane); hy = shfl(py, src_lane);
                            hit = true;
                            break;
                        }
                    }
                }
            }
        }
    }

    // ---- 5. where on the un-resampled path, and the cut (collision_avoidance.py:83-119, mpc_intersection.py:129-136)
    int cidx = -1, cutoff = ncf;
    double cxy_x = qnan, cxy_y = qnan;
    if (hit) {
        cidx = 0;                                                  // np.argmax of an all-False mask
        bool found = false;
        for (int c = 0; c < 2 && !found; ++c)
            for (int base = 0; base < n; base += 32) {
                const int i = base + lane;
                const bool h = i < n && within(hx, hy, pcc[(2 * c) * p.ncmax + aidx + i], pcc[(2 * c + 1) * p.ncmax + aidx + i], thr, thr2);
                const unsigned bal = __ballot_sync(MPCB_FULL, h);
                if (bal) { cidx = base + __ffs(bal) - 1; found = true; break; }
            }
        cxy_x = cx[aidx + cidx]; cxy_y = cy[aidx + cidx];
        int j = -1;
        for (int base = 0; base < ncf; base += 32) {
            const int i = base + lane;
            const bool h = i < ncf && within(cx[i], cy[i], cxy_x, cxy_y, 0.001, 0.001 * 0.001);
            const unsigned bal = __ballot_sync(MPCB_FULL, h);
            if (bal) { j = base + __ffs(bal) - 1; break; }
        }
        const int extra = 4 * (int)ceil(__ddiv_rn(p.radius, p.dl[route]));
        cutoff = max(aidx + 1, j - extra);
    }
    if (lane == 0) {
        p.agent_idx[b] = aidx;
        p.ncourse[b] = cutoff;
        if (p.collided) p.collided[b] = hit ? 1 : 0;
        if (p.coll_idx) p.coll_idx[b] = cidx;
        if (p.coll_xy) { p.coll_xy[2 * b] = cxy_x; p.coll_xy[2 * b + 1] = cxy_y; }
        if (p.n_res) p.n_res[b] = nres;
        if (p.flags) p.flags[b] = flags;
    }
}

}  // namespace mpcb
