// mpcb_astar.cu -- SURVEY.md 8f N2: the reference's motion-primitive A* planner, one warp per search.
//
// Replaces, per search, main/lib/motion_primitive_search_modified.py:54-58 (`MotionPrimitiveSearch.run`):
//   the search loop            main/lib/a_star.py:31-78 (heap order incl. tuple tie-breaks, lazy deletion, predecessor dict)
//   neighbour expansion        main/lib/motion_primitive_search_modified.py:101-149
//   collision check            main/lib/obstacles.py:157-176 (all half-planes <= 0 for any circle-centre point)
//   goal test / heuristic      main/lib/motion_primitive_search_modified.py:64-86, main/lib/obstacles.py:108-117
//   transform, angle wrap      main/lib/linalg.py:4-55, main/lib/maths.py:4-10
//   trajectory assembly        main/lib/motion_primitive_search_modified.py:151-165
// Integer outputs (primitive index sequence, path length) and the cost (a sequential sum of primitive lengths) are in the
// bit-exact class; node coordinates use a fixed operation order ((px*c + py*(-s)) + x, no FMA contraction), the one
// oracle/astar_oracle.py restates -- the reference itself multiplies through OpenBLAS, whose rounding is machine dependent.
//
// Data structures per search (global memory, sized by the caller):
//   open list   unsorted arrays in blocks of 256 entries with one running minimum of g+h per block; a pop is a warp-wide
//               arg-min by the reference's tuple order (g+h, g, node, predecessor) over the blocks whose minimum equals the
//               global one (almost always a single block), so its cost is n_open / 256 block minima + 256 entries instead
//               of n_open entries: the 10^4-expansion searches (10^5 entries) no longer set the launch time
//   closed set  arrays + an open-addressing hash table probed 32 slots at a time (exact float equality, like the dict)
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <string.h>

#include <string>
#include <vector>

#include "../../include/mpcb.h"

#define AS_FULL 0xffffffffu
#define AS_MAX_OBS 32
#define AS_MAX_HP 8
#define AS_MAX_MP 16
#define AS_MAX_CC 16
#define AS_OB 256                 /* open-list block: entries per running minimum */

namespace {

struct AstarParams {
    int B, S, P, NP, NCC, max_open, max_closed, hash_cap, max_path;
    const double *start, *goal, *goal_box, *allowed;      // [B,3] [B,3] [B,4] [B]
    const int* obs_set;                                   // [B]
    const double* hp;                                     // [S][32][8][3]
    const int* n_hp;                                      // [S][32]
    const int* n_obs;                                     // [S]
    const double* mp_points;                              // [P][NP][3]
    const int* mp_n;                                      // [P]
    const double* mp_len;                                 // [P]
    const double* mp_cc;                                  // [P][NCC][2]
    // workspace
    double *o_gh, *o_g, *o_x, *o_y, *o_t;                 // [B][max_open]
    int *o_pred, *o_mp;
    double* o_bmin;                                       // [B][max_open / AS_OB + 1] min g+h of the alive entries of a block
    double *c_x, *c_y, *c_t, *c_g;                        // [B][max_closed]
    int *c_pred, *c_mp;
    int* hash;                                            // [B][hash_cap], -1 = empty
    // outputs
    int *status, *n_path, *mp_seq, *n_expanded, *n_pushed, *n_traj;
    double *cost, *nodes, *traj;
};

struct AstarSmem {
    double hp[AS_MAX_OBS][AS_MAX_HP][3];
    int n_hp[AS_MAX_OBS];
    double cc[AS_MAX_MP * AS_MAX_CC][2];
    double last[AS_MAX_MP][3];
    double len[AS_MAX_MP];
};

__device__ __forceinline__ double inf_d() { return __longlong_as_double(0x7ff0000000000000LL); }

// Python float %: fmod, then the sign of the divisor (main/lib/maths.py:5)
__device__ __forceinline__ double normalize_angle(double theta) {
    const double tau = 6.283185307179586, pi = 3.141592653589793;
    double m = fmod(theta, tau);
    if (m != 0.0) { if (m < 0.0) m = __dadd_rn(m, tau); } else m = 0.0;
    if (m >= pi) m = __dsub_rn(m, tau);
    return m;
}

// sin / cos as a FIXED sequence of IEEE double operations (no libm, no FMA contraction): the same Cody-Waite reduction and
// polynomials, in the same order, as oracle/astar_oracle.py::sincos_fixed -- node coordinates are then bit-identical on the
// CPU and the GPU, and with them every tie the search breaks.  ~1 ulp; arguments here are |x| < 10.
__device__ __forceinline__ void sincos_fixed(double x, double* sn_out, double* cs_out) {
    const double P1 = 1.57079632673412561417e+00, P2 = 6.07710050630396597660e-11, P3 = 2.02226624871116645580e-21;
    const double S0 = -1.66666666666666324348e-01, S1 = 8.33333333332248946124e-03, S2 = -1.98412698298579493134e-04,
                 S3 = 2.75573137070700676789e-06, S4 = -2.50507602534068634195e-08, S5 = 1.58969099521155010221e-10;
    const double C0 = 4.16666666666666019037e-02, C1 = -1.38888888888741095749e-03, C2 = 2.48015872894767294178e-05,
                 C3 = -2.75573143513906633035e-07, C4 = 2.08757232129817482790e-09, C5 = -1.13596475577881948265e-11;
    const double k = rint(__dmul_rn(x, 6.36619772367581382433e-01));
    const double r = __dsub_rn(__dsub_rn(__dsub_rn(x, __dmul_rn(k, P1)), __dmul_rn(k, P2)), __dmul_rn(k, P3));
    const double z = __dmul_rn(r, r);
    auto poly = [&](double a0, double a1, double a2, double a3, double a4, double a5) {
        double v = __dadd_rn(a4, __dmul_rn(z, a5));
        v = __dadd_rn(a3, __dmul_rn(z, v)); v = __dadd_rn(a2, __dmul_rn(z, v));
        v = __dadd_rn(a1, __dmul_rn(z, v)); v = __dadd_rn(a0, __dmul_rn(z, v));
        return v;
    };
    const double sn = __dadd_rn(r, __dmul_rn(__dmul_rn(r, z), poly(S0, S1, S2, S3, S4, S5)));
    const double cs = __dadd_rn(__dsub_rn(1.0, __dmul_rn(0.5, z)), __dmul_rn(__dmul_rn(z, z), poly(C0, C1, C2, C3, C4, C5)));
    const int q = ((int)k) & 3;
    *sn_out = q == 0 ? sn : q == 1 ? cs : q == 2 ? -sn : -cs;
    *cs_out = q == 0 ? cs : q == 1 ? -sn : q == 2 ? -cs : sn;
}

__device__ __forceinline__ uint32_t hash3(double x, double y, double t) {
    // +0.0 canonicalises -0.0 (Python: 0.0 == -0.0 and they hash alike)
    uint64_t a = (uint64_t)__double_as_longlong(x + 0.0), b = (uint64_t)__double_as_longlong(y + 0.0),
             c = (uint64_t)__double_as_longlong(t + 0.0);
    uint64_t h = a * 0x9E3779B97F4A7C15ull;
    h ^= (b + 0x7F4A7C15ull) * 0xC2B2AE3D27D4EB4Full + (h << 6) + (h >> 2);
    h ^= (c + 0x165667B1ull) * 0x94D049BB133111EBull + (h << 6) + (h >> 2);
    h ^= h >> 29;
    return (uint32_t)(h ^ (h >> 32));
}

struct Search {
    const AstarParams& p;
    int lane;
    double *o_gh, *o_g, *o_x, *o_y, *o_t, *c_x, *c_y, *c_t, *c_g;
    int *o_pred, *o_mp, *c_pred, *c_mp, *hash;
    double* o_bmin;
    double sx, sy, st;                                   // start node (the predecessor of the first entry is itself)

    __device__ void pred_xyz(int e, double& x, double& y, double& t) const {
        const int pr = o_pred[e];
        if (pr < 0) { x = sx; y = sy; t = st; } else { x = c_x[pr]; y = c_y[pr]; t = c_t[pr]; }
    }
    // strict "entry a sorts before entry b" on the tail of the tuple (g, node, predecessor); equal g+h assumed
    __device__ bool less_tail(int a, int b) const {
        double u = o_g[a], v = o_g[b];
        if (u != v) return u < v;
        u = o_x[a]; v = o_x[b]; if (u != v) return u < v;
        u = o_y[a]; v = o_y[b]; if (u != v) return u < v;
        u = o_t[a]; v = o_t[b]; if (u != v) return u < v;
        double ax, ay, at, bx, by, bt;
        pred_xyz(a, ax, ay, at); pred_xyz(b, bx, by, bt);
        if (ax != bx) return ax < bx;
        if (ay != by) return ay < by;
        if (at != bt) return at < bt;
        return a < b;                                     // identical tuples: any order gives the same search
    }
    // heappop: the alive entry that is smallest in the reference's tuple order, or -1.  Two levels: the global minimum of
    // g+h from the block minima, then the full tuple comparison inside the block(s) that attain it.
    __device__ int pop(int n_open) const {
        const int nblk = (n_open + AS_OB - 1) / AS_OB;
        double gmin = inf_d();
        for (int k = lane; k < nblk; k += 32) gmin = fmin(gmin, o_bmin[k]);
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) gmin = fmin(gmin, __shfl_xor_sync(AS_FULL, gmin, d));
        if (!(gmin < inf_d())) return -1;
        int best = -1;
        for (int k0 = 0; k0 < nblk; k0 += 32) {
            const int k = k0 + lane;
            unsigned cand = __ballot_sync(AS_FULL, k < nblk && o_bmin[k] == gmin);
            while (cand) {
                const int blk = k0 + __ffs(cand) - 1;
                cand &= cand - 1;
                const int hi = min(n_open, (blk + 1) * AS_OB);
                for (int i = blk * AS_OB + lane; i < hi; i += 32)
                    if (o_gh[i] == gmin && (best < 0 || less_tail(i, best))) best = i;
            }
        }
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) {
            const int oi = __shfl_xor_sync(AS_FULL, best, d);
            if (oi >= 0 && (best < 0 || less_tail(oi, best))) best = oi;
        }
        return best;
    }
    // entry e left the queue: its block's running minimum is recomputed from the entries that are still alive
    __device__ void refresh_block(int e, int n_open) const {
        const int blk = e / AS_OB, hi = min(n_open, (blk + 1) * AS_OB);
        double m = inf_d();
        for (int i = blk * AS_OB + lane; i < hi; i += 32) m = fmin(m, o_gh[i]);
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) m = fmin(m, __shfl_xor_sync(AS_FULL, m, d));
        if (lane == 0) o_bmin[blk] = m;
    }
    // `node in pred_dict`: closed index or -1; *slot receives where an insertion would go
    __device__ int lookup(double x, double y, double t, int* slot) const {
        const int ngroups = p.hash_cap >> 5;
        int grp = (int)(hash3(x, y, t) & (uint32_t)(ngroups - 1));
        for (int probe = 0; probe < ngroups; ++probe) {
            const int s = (grp << 5) + lane;
            const int v = hash[s];
            const bool hit = v >= 0 && c_x[v] == x && c_y[v] == y && c_t[v] == t;
            const unsigned mh = __ballot_sync(AS_FULL, hit), me = __ballot_sync(AS_FULL, v < 0);
            if (mh) return __shfl_sync(AS_FULL, v, __ffs(mh) - 1);
            if (me) { *slot = (grp << 5) + __ffs(me) - 1; return -1; }
            grp = (grp + 1) & (ngroups - 1);
        }
        *slot = -1;
        return -1;
    }
};

__global__ void __launch_bounds__(32) astar_kernel(const AstarParams p) {
    __shared__ AstarSmem sm;
    const int lane = threadIdx.x, b = blockIdx.x;
    if (b >= p.B) return;
    // ---- stage the scenario ---------------------------------------------------------------------------------------
    const int set = p.obs_set[b];
    const int n_obs = min(p.n_obs[set], AS_MAX_OBS);
    for (int i = lane; i < AS_MAX_OBS * AS_MAX_HP * 3; i += 32) (&sm.hp[0][0][0])[i] = p.hp[(size_t)set * AS_MAX_OBS * AS_MAX_HP * 3 + i];
    if (lane < AS_MAX_OBS) sm.n_hp[lane] = p.n_hp[set * AS_MAX_OBS + lane];
    for (int i = lane; i < p.P * p.NCC; i += 32) { sm.cc[i][0] = p.mp_cc[2 * i]; sm.cc[i][1] = p.mp_cc[2 * i + 1]; }
    if (lane < p.P) {
        const double* l3 = p.mp_points + ((size_t)lane * p.NP + (p.mp_n[lane] - 1)) * 3;
        sm.last[lane][0] = l3[0]; sm.last[lane][1] = l3[1]; sm.last[lane][2] = l3[2];
        sm.len[lane] = p.mp_len[lane];
    }
    __syncwarp();
    const double gx = p.goal[3 * b], gy = p.goal[3 * b + 1], gth = p.goal[3 * b + 2];
    const double bx1 = p.goal_box[4 * b], by1 = p.goal_box[4 * b + 1], bx2 = p.goal_box[4 * b + 2], by2 = p.goal_box[4 * b + 3];
    const double allowed = p.allowed[b];
    Search S{p, lane,
             p.o_gh + (size_t)b * p.max_open, p.o_g + (size_t)b * p.max_open, p.o_x + (size_t)b * p.max_open,
             p.o_y + (size_t)b * p.max_open, p.o_t + (size_t)b * p.max_open,
             p.c_x + (size_t)b * p.max_closed, p.c_y + (size_t)b * p.max_closed, p.c_t + (size_t)b * p.max_closed,
             p.c_g + (size_t)b * p.max_closed,
             p.o_pred + (size_t)b * p.max_open, p.o_mp + (size_t)b * p.max_open,
             p.c_pred + (size_t)b * p.max_closed, p.c_mp + (size_t)b * p.max_closed, p.hash + (size_t)b * p.hash_cap,
             p.o_bmin + (size_t)b * (p.max_open / AS_OB + 1),
             p.start[3 * b], p.start[3 * b + 1], p.start[3 * b + 2]};
    for (int i = lane; i < p.hash_cap; i += 32) S.hash[i] = -1;
    if (lane == 0) {                                      // q = [(0, 0, start, start)]   (a_star.py:34)
        S.o_gh[0] = 0.0; S.o_g[0] = 0.0; S.o_x[0] = S.sx; S.o_y[0] = S.sy; S.o_t[0] = S.st; S.o_pred[0] = -1; S.o_mp[0] = -1;
        S.o_bmin[0] = 0.0;
    }
    __syncwarp();
    int n_open = 1, n_closed = 0, n_expanded = 0, status = MPCB_PLAN_NO_SOLUTION, goal_ci = -1;
    double goal_g = 0.0;

    while (true) {
        const int e = S.pop(n_open);
        if (e < 0) break;                                 // raise Exception("No solution found.")   (a_star.py:78)
        const double g = S.o_g[e], x = S.o_x[e], y = S.o_y[e], th = S.o_t[e];
        const int pred = S.o_pred[e], mp_in = S.o_mp[e];
        __syncwarp();
        if (lane == 0) S.o_gh[e] = inf_d();
        __syncwarp();
        S.refresh_block(e, n_open);
        int slot;
        int ci = S.lookup(x, y, th, &slot);
        if (ci >= 0 && g >= S.c_g[ci]) { __syncwarp(); continue; }      // seen before, not better   (a_star.py:45-49)
        ++n_expanded;
        if (ci < 0) {
            if (n_closed >= p.max_closed || slot < 0) { status = MPCB_PLAN_OVERFLOW; break; }
            ci = n_closed++;
            if (lane == 0) { S.c_x[ci] = x; S.c_y[ci] = y; S.c_t[ci] = th; S.hash[slot] = ci; }
        }
        if (lane == 0) { S.c_g[ci] = g; S.c_pred[ci] = pred; S.c_mp[ci] = mp_in; }     // pred_dict[node] = g, predecessor
        __syncwarp();
        // is_goal (motion_primitive_search_modified.py:64-69, obstacles.py:108-117)
        {
            const double dx = fmax(fmax(__dsub_rn(bx1, x), 0.0), __dsub_rn(x, bx2));
            const double dy = fmax(fmax(__dsub_rn(by1, y), 0.0), __dsub_rn(y, by2));
            const double dist = __dsqrt_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
            if (dist <= 1e-5 && fabs(__dsub_rn(th, gth)) <= allowed) { status = MPCB_PLAN_OK; goal_ci = ci; goal_g = g; break; }
        }
        // ---- neighbour expansion (motion_primitive_search_modified.py:101-149) ----------------------------------------
        double sn, cs;
        sincos_fixed(th, &sn, &cs);
        unsigned coll = 0;                                // bit m: primitive m collides
        const int npair = p.P * p.NCC;
        for (int q0 = 0; q0 < npair; q0 += 32) {
            const int q = q0 + lane;
            unsigned mine = 0;
            if (q < npair) {
                const double px = sm.cc[q][0], py = sm.cc[q][1];
                const double wx = __dadd_rn(__dadd_rn(__dmul_rn(px, cs), __dmul_rn(py, -sn)), x);
                const double wy = __dadd_rn(__dadd_rn(__dmul_rn(px, sn), __dmul_rn(py, cs)), y);
                bool hit = false;
                for (int o = 0; o < n_obs && !hit; ++o) {
                    bool inside = true;
                    const int nh = sm.n_hp[o];
                    for (int k = 0; k < nh; ++k) {
                        const double v = __dadd_rn(__dadd_rn(__dmul_rn(sm.hp[o][k][0], wx), __dmul_rn(sm.hp[o][k][1], wy)), sm.hp[o][k][2]);
                        inside = inside && (v <= 0.0);
                    }
                    hit = inside;
                }
                if (hit) mine = 1u << (q / p.NCC);
            }
            coll |= __reduce_or_sync(AS_FULL, mine);
        }
        for (int m = 0; m < p.P; ++m) {
            if (coll & (1u << m)) continue;
            const double lx = sm.last[m][0], ly = sm.last[m][1];
            const double nx = __dadd_rn(__dadd_rn(__dmul_rn(lx, cs), __dmul_rn(ly, -sn)), x);
            const double ny = __dadd_rn(__dadd_rn(__dmul_rn(lx, sn), __dmul_rn(ly, cs)), y);
            const double nt = normalize_angle(__dadd_rn(sm.last[m][2], th));
            const double ng = __dadd_rn(g, sm.len[m]);
            int dummy;
            const int cj = S.lookup(nx, ny, nt, &dummy);
            if (cj >= 0 && !(ng < S.c_g[cj])) continue;                  // a_star.py:71
            if (n_open >= p.max_open) { status = MPCB_PLAN_OVERFLOW; break; }
            if (lane == 0) {
                const double ddx = __dsub_rn(nx, gx), ddy = __dsub_rn(ny, gy);
                const double dxy = __dsqrt_rn(__dadd_rn(__dmul_rn(ddx, ddx), __dmul_rn(ddy, ddy)));
                const double ad = fabs(__dsub_rn(nt, gth));
                const double dth = fmin(ad, __dsub_rn(ad, allowed / 2));
                const double h = __dadd_rn(dxy, __dmul_rn(2.7, dth));      // :71-86
                const double gh = __dadd_rn(ng, h);
                S.o_gh[n_open] = gh; S.o_g[n_open] = ng; S.o_x[n_open] = nx; S.o_y[n_open] = ny; S.o_t[n_open] = nt;
                S.o_pred[n_open] = ci; S.o_mp[n_open] = m;
                const int blk = n_open / AS_OB;
                S.o_bmin[blk] = (n_open % AS_OB == 0) ? gh : fmin(S.o_bmin[blk], gh);
            }
            ++n_open;
        }
        __syncwarp();
        if (status == MPCB_PLAN_OVERFLOW) break;
    }

    // ---- path reconstruction (a_star.py:58-66) and trajectory (motion_primitive_search_modified.py:151-165) ----------
    int n_path = 0, n_traj = 0;
    int* seq = p.mp_seq + (size_t)b * p.max_path;
    double* nodes = p.nodes + (size_t)b * p.max_path * 3;
    if (status == MPCB_PLAN_OK) {
        int len = 1;
        for (int c = goal_ci; S.c_pred[c] >= 0; c = S.c_pred[c]) ++len;
        if (len > p.max_path) status = MPCB_PLAN_OVERFLOW;
        else {
            n_path = len;
            if (lane == 0) {
                int c = goal_ci;
                for (int i = len - 1; i >= 0; --i) {
                    nodes[3 * i] = S.c_x[c]; nodes[3 * i + 1] = S.c_y[c]; nodes[3 * i + 2] = S.c_t[c];
                    if (i > 0) seq[i - 1] = S.c_mp[c];
                    c = S.c_pred[c];
                }
                for (int i = len - 1; i < p.max_path; ++i) seq[i] = -1;
            }
            __syncwarp();
            if (p.traj) {
                double* tr = p.traj + (size_t)b * (p.max_path - 1) * (p.NP - 1) * 3;
                for (int i = 0; i + 1 < len; ++i) {
                    const int m = seq[i];
                    const double x = nodes[3 * i], y = nodes[3 * i + 1], th = nodes[3 * i + 2];
                    double sn, cs;
                    sincos_fixed(th, &sn, &cs);
                    const int np1 = p.mp_n[m] - 1;
                    for (int k = lane; k < np1; k += 32) {
                        const double* q3 = p.mp_points + ((size_t)m * p.NP + k) * 3;
                        double* o3 = tr + (size_t)(n_traj + k) * 3;
                        o3[0] = __dadd_rn(__dadd_rn(__dmul_rn(q3[0], cs), __dmul_rn(q3[1], -sn)), x);
                        o3[1] = __dadd_rn(__dadd_rn(__dmul_rn(q3[0], sn), __dmul_rn(q3[1], cs)), y);
                        o3[2] = __dadd_rn(q3[2], th);
                    }
                    n_traj += np1;
                }
            }
        }
    }
    if (lane == 0) {
        p.status[b] = status;
        p.cost[b] = status == MPCB_PLAN_OK ? goal_g : __longlong_as_double(0x7ff8000000000000LL);
        p.n_path[b] = n_path;
        p.n_expanded[b] = n_expanded;
        p.n_pushed[b] = n_open;
        if (p.n_traj) p.n_traj[b] = n_traj;
        if (status != MPCB_PLAN_OK) for (int i = 0; i < p.max_path; ++i) seq[i] = -1;
    }
}

thread_local std::string g_plan_error;

}  // namespace

extern "C" {

const char* mpcb_plan_last_error(void) { return g_plan_error.c_str(); }

int mpcb_plan_routes_host(int device, int B, const double* start, const double* goal, const double* goal_box,
                          const double* allowed, const int* obs_set, int S, const double* halfplanes, const int* n_halfplanes,
                          const int* n_obstacles, int P, int NP, const double* mp_points, const int* mp_n_points,
                          const double* mp_length, int NCC, const double* mp_cc_points, int max_open, int max_closed,
                          int max_path, int* status, double* cost, int* n_path, int* mp_seq, double* nodes, int* n_expanded,
                          int* n_pushed, double* traj, int* n_traj) {
    if (B <= 0 || !start || !goal || !goal_box || !allowed || !obs_set || S <= 0 || !halfplanes || !n_halfplanes || !n_obstacles ||
        P <= 0 || P > AS_MAX_MP || NP < 2 || !mp_points || !mp_n_points || !mp_length || NCC <= 0 || NCC > AS_MAX_CC || !mp_cc_points ||
        max_open < 16 || max_closed < 16 || max_path < 2 || !status || !cost || !n_path || !mp_seq || !nodes) {
        g_plan_error = "mpcb_plan_routes_host: bad arguments (P <= 16 primitives, NCC <= 16 collision points each)";
        return MPCB_EINVAL;
    }
    for (int b = 0; b < B; ++b) if (obs_set[b] < 0 || obs_set[b] >= S) { g_plan_error = "mpcb_plan_routes_host: obs_set out of range"; return MPCB_EINVAL; }
    for (int s = 0; s < S; ++s) {
        if (n_obstacles[s] < 0 || n_obstacles[s] > AS_MAX_OBS) { g_plan_error = "mpcb_plan_routes_host: at most 32 obstacles per set"; return MPCB_EINVAL; }
        for (int o = 0; o < n_obstacles[s]; ++o)
            if (n_halfplanes[s * AS_MAX_OBS + o] < 1 || n_halfplanes[s * AS_MAX_OBS + o] > AS_MAX_HP) { g_plan_error = "mpcb_plan_routes_host: 1..8 half-planes per obstacle"; return MPCB_EINVAL; }
    }
    for (int m = 0; m < P; ++m) if (mp_n_points[m] < 2 || mp_n_points[m] > NP) { g_plan_error = "mpcb_plan_routes_host: mp_n_points out of range"; return MPCB_EINVAL; }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0 || device < 0 || device >= ndev) { g_plan_error = "mpcb_plan_routes_host: no usable CUDA device (there is no CPU fallback)"; return MPCB_ENODEV; }
    if (cudaSetDevice(device) != cudaSuccess) { g_plan_error = "cudaSetDevice failed"; return MPCB_ECUDA; }
    int hash_cap = 64;
    while (hash_cap < 2 * max_closed) hash_cap <<= 1;
    const size_t nB = (size_t)B;
    size_t off = 0;
    auto put = [&](size_t bytes) { size_t at = off; off = (off + bytes + 255) & ~(size_t)255; return at; };
    const size_t i_start = put(24 * nB), i_goal = put(24 * nB), i_box = put(32 * nB), i_allowed = put(8 * nB), i_set = put(4 * nB),
                 i_hp = put(8 * 3 * AS_MAX_HP * AS_MAX_OBS * (size_t)S), i_nhp = put(4 * AS_MAX_OBS * (size_t)S), i_nobs = put(4 * (size_t)S),
                 i_pts = put(8 * 3 * (size_t)P * NP), i_mpn = put(4 * (size_t)P), i_len = put(8 * (size_t)P), i_cc = put(16 * (size_t)P * NCC),
                 w_gh = put(8 * nB * max_open), w_g = put(8 * nB * max_open), w_x = put(8 * nB * max_open), w_y = put(8 * nB * max_open),
                 w_t = put(8 * nB * max_open), w_pred = put(4 * nB * max_open), w_mp = put(4 * nB * max_open), w_bmin = put(8 * nB * ((size_t)max_open / AS_OB + 1)),
                 c_x = put(8 * nB * max_closed), c_y = put(8 * nB * max_closed), c_t = put(8 * nB * max_closed), c_g = put(8 * nB * max_closed),
                 c_pred = put(4 * nB * max_closed), c_mp = put(4 * nB * max_closed), w_hash = put(4 * nB * hash_cap),
                 o_status = put(4 * nB), o_cost = put(8 * nB), o_npath = put(4 * nB), o_seq = put(4 * nB * max_path),
                 o_nodes = put(24 * nB * max_path), o_nexp = put(4 * nB), o_npush = put(4 * nB), o_ntraj = put(4 * nB),
                 o_traj = put(traj ? 24 * nB * (size_t)(max_path - 1) * (NP - 1) : 0);
    char* d = nullptr;
    if (cudaMalloc(&d, off) != cudaSuccess) { g_plan_error = "mpcb_plan_routes_host: cudaMalloc of the search workspace failed (lower max_open / max_closed or the batch)"; return MPCB_ENOMEM; }
    int rc = MPCB_OK;
    auto up = [&](size_t at, const void* src, size_t bytes) { if (cudaMemcpy(d + at, src, bytes, cudaMemcpyHostToDevice) != cudaSuccess) rc = MPCB_ECUDA; };
    up(i_start, start, 24 * nB); up(i_goal, goal, 24 * nB); up(i_box, goal_box, 32 * nB); up(i_allowed, allowed, 8 * nB); up(i_set, obs_set, 4 * nB);
    up(i_hp, halfplanes, 8 * 3 * AS_MAX_HP * AS_MAX_OBS * (size_t)S); up(i_nhp, n_halfplanes, 4 * AS_MAX_OBS * (size_t)S); up(i_nobs, n_obstacles, 4 * (size_t)S);
    up(i_pts, mp_points, 8 * 3 * (size_t)P * NP); up(i_mpn, mp_n_points, 4 * (size_t)P); up(i_len, mp_length, 8 * (size_t)P); up(i_cc, mp_cc_points, 16 * (size_t)P * NCC);
    AstarParams p;
    memset(&p, 0, sizeof(p));
    p.B = B; p.S = S; p.P = P; p.NP = NP; p.NCC = NCC; p.max_open = max_open; p.max_closed = max_closed; p.hash_cap = hash_cap; p.max_path = max_path;
    p.start = (double*)(d + i_start); p.goal = (double*)(d + i_goal); p.goal_box = (double*)(d + i_box); p.allowed = (double*)(d + i_allowed);
    p.obs_set = (int*)(d + i_set); p.hp = (double*)(d + i_hp); p.n_hp = (int*)(d + i_nhp); p.n_obs = (int*)(d + i_nobs);
    p.mp_points = (double*)(d + i_pts); p.mp_n = (int*)(d + i_mpn); p.mp_len = (double*)(d + i_len); p.mp_cc = (double*)(d + i_cc);
    p.o_gh = (double*)(d + w_gh); p.o_g = (double*)(d + w_g); p.o_x = (double*)(d + w_x); p.o_y = (double*)(d + w_y); p.o_t = (double*)(d + w_t);
    p.o_pred = (int*)(d + w_pred); p.o_mp = (int*)(d + w_mp); p.o_bmin = (double*)(d + w_bmin);
    p.c_x = (double*)(d + c_x); p.c_y = (double*)(d + c_y); p.c_t = (double*)(d + c_t); p.c_g = (double*)(d + c_g);
    p.c_pred = (int*)(d + c_pred); p.c_mp = (int*)(d + c_mp); p.hash = (int*)(d + w_hash);
    p.status = (int*)(d + o_status); p.cost = (double*)(d + o_cost); p.n_path = (int*)(d + o_npath); p.mp_seq = (int*)(d + o_seq);
    p.nodes = (double*)(d + o_nodes); p.n_expanded = (int*)(d + o_nexp); p.n_pushed = (int*)(d + o_npush); p.n_traj = (int*)(d + o_ntraj);
    p.traj = traj ? (double*)(d + o_traj) : nullptr;
    if (rc == MPCB_OK) {
        astar_kernel<<<B, 32>>>(p);
        if (cudaGetLastError() != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) { g_plan_error = "mpcb_plan_routes_host: kernel failed"; rc = MPCB_ECUDA; }
    }
    auto down = [&](void* dst, size_t at, size_t bytes) { if (dst && rc == MPCB_OK && cudaMemcpy(dst, d + at, bytes, cudaMemcpyDeviceToHost) != cudaSuccess) rc = MPCB_ECUDA; };
    down(status, o_status, 4 * nB); down(cost, o_cost, 8 * nB); down(n_path, o_npath, 4 * nB); down(mp_seq, o_seq, 4 * nB * max_path);
    down(nodes, o_nodes, 24 * nB * max_path); down(n_expanded, o_nexp, 4 * nB); down(n_pushed, o_npush, 4 * nB);
    down(n_traj, o_ntraj, 4 * nB);
    if (traj) down(traj, o_traj, 24 * nB * (size_t)(max_path - 1) * (NP - 1));
    cudaFree(d);
    if (rc == MPCB_ECUDA && g_plan_error.empty()) g_plan_error = "mpcb_plan_routes_host: CUDA copy failed";
    return rc;
}

}  // extern "C"
