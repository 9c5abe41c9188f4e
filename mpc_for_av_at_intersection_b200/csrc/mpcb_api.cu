// mpcb_api.cu -- C ABI of libmpcb.so (see include/mpcb.h) and the small A8 kernels.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/mpcb.h"
#include "mpcb_step.cuh"
#include "mpcb_step_half.cuh"
#include "mpcb_step_long.cuh"
#include "mpcb_coupling.cuh"

namespace {

thread_local std::string g_create_error;

struct HostStage {
    void* pinned = nullptr;
    void* dev = nullptr;
    size_t bytes = 0;
};

}  // namespace

struct mpcb_handle {
    mpcb_config cfg;
    int device = 0;
    int max_batch = 0;
    int R = 0, ncmax = 0;
    double* courses_dev = nullptr;
    double* dl_dev = nullptr;
    long long launches = 0;
    std::string err;
    // staging for the *_host entry points
    HostStage stage;
    cudaStream_t own_stream = nullptr;
    int smem_optin = 0;
    int num_sms = 148;
    cudaStream_t last_stream = nullptr;   // stream of the last asynchronous session call (fetch/store wait on it)
    bool last_stream_set = false;
    int variant = 0;             // kernel tuning variant (MPCB_VARIANT, only in MPCB_TUNE_VARIANTS builds)
    int* len_dev = nullptr;
    double* pathcc_dev = nullptr;    // [R][2 circles][x|y][ncmax]
    double* ramp_dev = nullptr;      // [ncmax] np.cumsum(zeros + MAX_ACCEL): the ego prediction's acceleration ramp
    std::vector<int> len_host;
    std::vector<double> goal_host;   // [R,2] last point of each full course (mpc.py:263)
    // resident session: the [B,...] mirror of the reference's MPC object state lives on the device
    struct Session {
        int B = 0;
        char* slab = nullptr;
        char* pin = nullptr;        // pinned: state in | di_ai out | status out
        double *state = nullptr, *L = nullptr, *oa = nullptr, *od = nullptr, *traj = nullptr, *xref = nullptr,
               *di_ai = nullptr, *xdev = nullptr, *goal = nullptr;
        int *route = nullptr, *ncourse = nullptr, *target = nullptr, *warm = nullptr, *status = nullptr,
            *iters = nullptr, *done = nullptr, *agent_idx = nullptr, *collided = nullptr, *coll_idx = nullptr,
            *n_script = nullptr;
        double *script = nullptr, *obs = nullptr;      // scripted cross traffic [B,8,8] and its last get() rows [B,8,6]
        double* pw = nullptr;                          // per-problem cost weights [B,12]
        double* lam_ws = nullptr;                      // interior-point multipliers of the last solve [B,8,T] (warm start)
        int* vcut = nullptr;                           // mpc_with_speed cutoff_idx [B]
        int *lpt_order = nullptr, *lpt_rank = nullptr, *lpt_hist = nullptr;   // longest-first launch order (mpcb_step.cuh)
        bool hasL = false, has_script = false, has_pw = false, has_vcut = false;
        std::vector<int> route_host;                   // host copy of the routes: the host-array entry points validate indices against it
        size_t pin_state = 0, pin_diai = 0, pin_status = 0, pin_total = 0;
        char* pin_all = nullptr; size_t pin_all_bytes = 0;     // staging of mpcb_session_step_all_host (allocated on first use)
    } ses;
};

#define CUDA_TRY(h, expr)                                                                      \
    do {                                                                                       \
        cudaError_t e_ = (expr);                                                               \
        if (e_ != cudaSuccess) {                                                               \
            (h)->err = std::string(#expr) + ": " + cudaGetErrorString(e_);                     \
            return MPCB_ECUDA;                                                                 \
        }                                                                                      \
    } while (0)

// ---------------------------------------------------------------------------------------------------
// A8 kernels: plant step and goal test (one thread per problem; HBM-bound, 32-40 B per problem)
// ---------------------------------------------------------------------------------------------------
// main/lib/simulation.py:35-47 + main/bicycle/main.py:28-41; explicit rn ops mirror numpy's unfused arithmetic
__global__ void plant_step_kernel(int B, double* __restrict__ state, const double* __restrict__ di_ai,
                                  const double* __restrict__ Lp, const int* __restrict__ active, double L0,
                                  double dt, double max_steer, double max_speed, double min_speed) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    if (active && !active[b]) return;
    double4 st = reinterpret_cast<double4*>(state)[b];       // x y yaw v
    const double2 u = reinterpret_cast<const double2*>(di_ai)[b];   // steer, accel
    const double L = Lp ? Lp[b] : L0;
    const double delta = fmax(fmin(u.x, max_steer), -max_steer);
    double sn, cs;
    sincos(st.z, &sn, &cs);
    const double xd = __dmul_rn(st.w, cs), yd = __dmul_rn(st.w, sn);
    const double td = __dmul_rn(__ddiv_rn(st.w, L), tan(delta));
    st.x = __dadd_rn(st.x, __dmul_rn(xd, dt));
    st.y = __dadd_rn(st.y, __dmul_rn(yd, dt));
    st.z = __dadd_rn(st.z, __dmul_rn(td, dt));
    double v = __dadd_rn(st.w, __dmul_rn(u.y, dt));
    st.w = fmax(fmin(v, max_speed), min_speed);
    reinterpret_cast<double4*>(state)[b] = st;
}

// main/lib/mpc.py:310-326
__global__ void is_goal_kernel(int B, const double* __restrict__ state, const double* __restrict__ goal,
                               const int* __restrict__ target_ind, const int* __restrict__ ncourse,
                               int* __restrict__ out, double goal_dis, double stop_speed) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const double4 st = reinterpret_cast<const double4*>(state)[b];
    const double2 g = reinterpret_cast<const double2*>(goal)[b];
    const double d = hypot(st.x - g.x, st.y - g.y);
    bool isgoal = d <= goal_dis;
    if (abs(target_ind[b] - ncourse[b]) >= 5) isgoal = false;
    out[b] = (isgoal && fabs(st.w) <= stop_speed) ? 1 : 0;
}

// FP64 FMA peak: 8 independent chains per thread, enough warps to fill every SM
__global__ void fp64_peak_kernel(double* out, int iters, double seed) {
    double a0 = seed, a1 = seed + 1, a2 = seed + 2, a3 = seed + 3, a4 = seed + 4, a5 = seed + 5, a6 = seed + 6, a7 = seed + 7;
    const double m = 1.0000001, c = 1e-9;
    for (int i = 0; i < iters; ++i) {
        a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
        a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}

// ---------------------------------------------------------------------------------------------------
// step kernel dispatch
// ---------------------------------------------------------------------------------------------------
namespace {

constexpr int kHorizons[] = {10, 13, 20, 40, 80};      // <= 31: one warp per problem; longer: one CTA per problem
constexpr int kLongThreads = 128;

template <int T, int WARPS, int MINB, bool VAR = false>
int launch_step(mpcb_handle* h, const StepParams& p, cudaStream_t st) {
    const size_t smem = sizeof(mpcb::WarpSmem<T>) * WARPS;
    void (*kern)(const StepParams);
    if constexpr (VAR) kern = mpcb::mpc_step_var_kernel<T, WARPS, MINB>;
    else if constexpr (MINB > 32) kern = mpcb::mpc_step_kernel_r<T, WARPS, MINB>;
    else kern = mpcb::mpc_step_kernel<T, WARPS, MINB>;
    static thread_local int configured_dev = -1;
    if (configured_dev != h->device) {
        CUDA_TRY(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        CUDA_TRY(h, cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
        configured_dev = h->device;
#ifdef MPCB_TUNE_VARIANTS
        if (getenv("MPCB_SHOW_OCCUPANCY")) {
            int nb = 0;
            cudaFuncAttributes fa;
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kern, WARPS * 32, smem);
            cudaFuncGetAttributes(&fa, kern);
            int nb0 = 0, nbh = 0, smsm = 0, regsm = 0;
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb0, kern, WARPS * 32, 0);
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nbh, kern, WARPS * 32, smem / 2);
            cudaDeviceGetAttribute(&smsm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, h->device);
            cudaDeviceGetAttribute(&regsm, cudaDevAttrMaxRegistersPerMultiprocessor, h->device);
            fprintf(stderr, "mpcb: smem/SM=%d regs/SM=%d CTAs/SM with smem 0: %d, smem/2: %d\n", smsm, regsm, nb0, nbh);
            fprintf(stderr, "mpcb: T=%d warps/CTA=%d bound=%d regs=%d smem/CTA=%zu local=%zu resident CTAs/SM=%d\n", T, WARPS, MINB,
                    fa.numRegs, smem, (size_t)fa.localSizeBytes, nb);
        }
#endif
    }
    const int grid = (p.B + WARPS - 1) / WARPS;
    kern<<<grid, WARPS * 32, smem, st>>>(p);
    CUDA_TRY(h, cudaGetLastError());
    h->launches++;
    return MPCB_OK;
}

// Two problems per warp (mpcb_step_half.cuh): T + 1 <= 16.  MPCB_HALF_WARP=0 in the environment falls back to the lane-per-stage
// kernel (A/B runs; read once).
template <int T, int WARPS, int MINB, bool VAR = false>
int launch_step_half(mpcb_handle* h, const StepParams& p, cudaStream_t st) {
    const size_t smem = 2 * sizeof(mpcb::WarpSmem<T>) * WARPS;
    void (*kern)(const StepParams);
    if constexpr (VAR) kern = mpcb::mpc_step_half_var_kernel<T, WARPS, MINB>;
    else kern = mpcb::mpc_step_half_kernel<T, WARPS, MINB>;
    static thread_local int configured_dev = -1;
    if (configured_dev != h->device) {
        CUDA_TRY(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        CUDA_TRY(h, cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
        configured_dev = h->device;
    }
    const int warps = (p.B + 1) / 2;
    const int grid = (warps + WARPS - 1) / WARPS;
    kern<<<grid, WARPS * 32, smem, st>>>(p);
    CUDA_TRY(h, cudaGetLastError());
    h->launches++;
    return MPCB_OK;
}
bool half_warp_enabled() {
    static const bool on = [] { const char* e = getenv("MPCB_HALF_WARP"); return !(e && e[0] == '0'); }();
    return on;
}

int dispatch_step(mpcb_handle* h, const StepParams& p, cudaStream_t st);

// long horizons (mpcb_step_long.cuh): one CTA per problem, system matrix in shared memory
template <int NT, int MINB = 1, bool JERK = false>
int launch_step_long_nt(mpcb_handle* h, const StepParams& p, cudaStream_t st) {
    const int T = h->cfg.T;
    if (!JERK && (T + 1 > NT || T % 4 != 0)) { h->err = "mpcb: the CTA-per-problem kernel needs T + 1 <= threads and T a multiple of 4 (8x8 tiles)"; return MPCB_EINVAL; }
    if (JERK && (int)mpcb::LongSmem::unknowns(T, true) > NT) { h->err = "mpcb: the jerk kernel needs 2 T + 1 <= 64 unknowns"; return MPCB_EINVAL; }
    const size_t smem = mpcb::LongSmem::bytes(T, NT, JERK);
    auto kern = mpcb::mpc_step_long_kernel<NT, MINB, JERK>;
    static thread_local int configured_dev = -1;
    static thread_local size_t configured_smem = 0;
    if (configured_dev != h->device || configured_smem < smem) {
        CUDA_TRY(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        CUDA_TRY(h, cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
        configured_dev = h->device; configured_smem = smem;
    }
    kern<<<p.B, NT, smem, st>>>(p, T);
    CUDA_TRY(h, cudaGetLastError());
    h->launches++;
    return MPCB_OK;
}
// Threads per problem (rows of the system matrix go to threads i % NT).  T = 80: 146 KB of shared memory = one CTA per SM
// whatever its size; 8 warps sharing the tile updates and the row phases: 111 K solves/s, 4 warps: 97 K (B = 4096).  T = 40:
// 48 KB per problem; 64 threads (4 CTAs per SM, every barrier stall of one problem hidden by three others): 766 K, 128
// threads (2 CTAs per SM at 250 registers): 531 K; 128 threads capped at 168 / 128 registers (3 / 4 CTAs, spills): 512 K / 593 K.
int launch_step_long(mpcb_handle* h, const StepParams& p, cudaStream_t st) {
    // threads per problem (MPCB_LONG_THREADS overrides, tuning only): rows of the system matrix go to threads i % NT
    static const int nt_env = getenv("MPCB_LONG_THREADS") ? atoi(getenv("MPCB_LONG_THREADS")) : 0;
    const int nthreads = nt_env ? nt_env : (h->cfg.T > 40 ? 256 : 64);
    if (nthreads == 256) return launch_step_long_nt<256>(h, p, st);
    if (nthreads == 64) return launch_step_long_nt<64, 4>(h, p, st);
    return launch_step_long_nt<kLongThreads>(h, p, st);      // 128 threads: measured and rejected at T = 40 (profiles/r02_NOTES.md)
}

#ifdef MPCB_TRACE
double* g_trace_dev = nullptr;      // development builds only (scripts/trace_failures.py): per-iteration log [B][128][8]
#endif

StepParams make_params(const mpcb_handle* h) {
    const mpcb_config& c = h->cfg;
    StepParams p;
    memset(&p, 0, sizeof(p));
#ifdef MPCB_TRACE
    p.trace = g_trace_dev;
#endif
    p.R = h->R; p.ncmax = h->ncmax; p.courses = h->courses_dev; p.dl = h->dl_dev;
    p.w_perp = c.w_perp; p.w_para = c.w_para; p.R0 = c.R[0]; p.R1 = c.R[1]; p.Rd0 = c.Rd[0]; p.Rd1 = c.Rd[1];
    p.Qv = c.Q_v_yaw[0]; p.Qyaw = c.Q_v_yaw[1];
    p.Qf0 = c.Qf[0] * c.T; p.Qf1 = c.Qf[1] * c.T; p.Qf2 = c.Qf[2] * c.T; p.Qf3 = c.Qf[3] * c.T;   // mpc.py:25
    p.dsteer_dt = (c.max_dsteer_deg * (3.141592653589793 / 180.0)) * c.dt;                        // mpc.py:34,184
    p.max_accel = c.max_accel; p.max_decel = c.max_decel; p.max_steer = c.max_steer;
    p.max_speed = c.max_speed; p.min_speed = c.min_speed; p.r_end = c.r_end; p.min_ref_speed = c.min_ref_speed;
    p.dt = c.dt; p.wheelbase = c.wheelbase; p.max_iter = c.max_iter; p.ipm_max_iter = c.ipm_max_iter;
    p.tol_mu = c.ipm_tol_mu; p.tol_p = c.ipm_tol_p; p.tol_d = c.ipm_tol_d;
    p.goal_dis = c.goal_dis; p.stop_speed = c.stop_speed;
    p.ref_speed = c.ref_speed; p.jerk_weight = c.jerk_weight;
    return p;
}

}  // namespace

extern "C" {

int mpcb_version(void) { return MPCB_VERSION; }

void mpcb_default_config(mpcb_config* c) {
    memset(c, 0, sizeof(*c));
    c->T = 13; c->max_iter = 1; c->w_perp = 20.0; c->w_para = 1.0;
    c->R[0] = 0.01; c->R[1] = 0.01; c->Rd[0] = 0.01; c->Rd[1] = 1.0;
    c->Q_v_yaw[0] = 0.0; c->Q_v_yaw[1] = 0.5;
    c->Qf[0] = 1.0; c->Qf[1] = 1.0; c->Qf[2] = 0.0; c->Qf[3] = 0.5;
    c->goal_dis = 1.5; c->stop_speed = 0.1389; c->max_dsteer_deg = 30.0; c->max_accel = 2.0; c->max_decel = -10.0;
    c->max_steer = 45.0 * (3.141592653589793 / 180.0);      // np.deg2rad(45.0)
    c->max_speed = 30.0 / 3.6; c->min_speed = -5.0; c->r_end = 10.0; c->min_ref_speed = 10 / 3.6;
    c->dt = 0.2; c->wheelbase = 2.86;
    c->ipm_max_iter = 100; c->ipm_tol_mu = 1e-12; c->ipm_tol_p = 1e-10; c->ipm_tol_d = 1e-9;
    c->car_width = 2.0; c->car_length = 2.86 + 0.64;      // BicycleModelDimensions (car_dimensions.py:82-90)
    c->coupling_horizon = 7.0; c->frame_window = 20;      // mpc_intersection.py:81-82
    c->jerk_model = 0; c->jerk_weight = 1.0;              // mpc_jerk.py:30
}

int mpcb_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return MPCB_ENODEV;
    return n;
}

int mpcb_supported_horizons(int* out, int cap) {
    const int n = (int)(sizeof(kHorizons) / sizeof(kHorizons[0]));
    for (int i = 0; i < n && i < cap; ++i) out[i] = kHorizons[i];
    return n;
}

const char* mpcb_last_error(const mpcb_handle* h) { return h ? h->err.c_str() : g_create_error.c_str(); }

int mpcb_create(const mpcb_config* cfg, int max_batch, int device, mpcb_handle** out) {
    if (!cfg || !out || max_batch <= 0) { g_create_error = "mpcb_create: bad arguments"; return MPCB_EINVAL; }
    bool okT = false;
    for (int t : kHorizons) okT = okT || (t == cfg->T);
    if (!okT) { g_create_error = "mpcb_create: horizon T=" + std::to_string(cfg->T) + " has no kernel in this build"; return MPCB_EINVAL; }
    if (cfg->max_iter < 1 || cfg->dt <= 0) { g_create_error = "mpcb_create: bad config"; return MPCB_EINVAL; }
    if (cfg->jerk_model && cfg->T > 28) { g_create_error = "mpcb_create: the jerk model (main/lib/mpc_jerk.py) is built for T <= 28"; return MPCB_EINVAL; }
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0 || device < 0 || device >= n) {
        g_create_error = "mpcb_create: no usable CUDA device (this library has no CPU fallback)";
        return MPCB_ENODEV;
    }
    mpcb_handle* h = new mpcb_handle();
    h->cfg = *cfg; h->device = device; h->max_batch = max_batch;
    if (cudaSetDevice(device) != cudaSuccess) { g_create_error = "cudaSetDevice failed"; delete h; return MPCB_ECUDA; }
    cudaDeviceGetAttribute(&h->smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device);
    cudaDeviceGetAttribute(&h->num_sms, cudaDevAttrMultiProcessorCount, device);
    if (const char* v = getenv("MPCB_VARIANT")) h->variant = atoi(v);
    if (cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking) != cudaSuccess) {
        g_create_error = "cudaStreamCreate failed"; delete h; return MPCB_ECUDA;
    }
    *out = h;
    return MPCB_OK;
}

void mpcb_destroy(mpcb_handle* h) {
    if (!h) return;
    cudaSetDevice(h->device);
    if (h->courses_dev) cudaFree(h->courses_dev);
    if (h->dl_dev) cudaFree(h->dl_dev);
    if (h->len_dev) cudaFree(h->len_dev);
    if (h->pathcc_dev) cudaFree(h->pathcc_dev);
    if (h->ramp_dev) cudaFree(h->ramp_dev);
    if (h->ses.slab) cudaFree(h->ses.slab);
    if (h->ses.pin) cudaFreeHost(h->ses.pin);
    if (h->ses.pin_all) cudaFreeHost(h->ses.pin_all);
    if (h->stage.pinned) cudaFreeHost(h->stage.pinned);
    if (h->stage.dev) cudaFree(h->stage.dev);
    if (h->own_stream) cudaStreamDestroy(h->own_stream);
    delete h;
}

void mpcb_smooth_yaw(double* yaw, int n) {
    const double pi = 3.141592653589793;
    for (int i = 0; i + 1 < n; ++i) {
        double dyaw = yaw[i + 1] - yaw[i];
        while (dyaw >= pi / 2.0) { yaw[i + 1] -= pi * 2.0; dyaw = yaw[i + 1] - yaw[i]; }
        while (dyaw <= -pi / 2.0) { yaw[i + 1] += pi * 2.0; dyaw = yaw[i + 1] - yaw[i]; }
    }
}

int mpcb_set_courses_host(mpcb_handle* h, const double* courses_host, const int* len_host, const double* dl_host,
                          int R, int ncmax) {
    if (!h || !courses_host || !len_host || !dl_host || R <= 0 || ncmax <= 0) { if (h) h->err = "mpcb_set_courses_host: bad arguments"; return MPCB_EINVAL; }
    for (int r = 0; r < R; ++r)
        if (len_host[r] < 1 || len_host[r] > ncmax) { h->err = "mpcb_set_courses_host: course length out of range"; return MPCB_EINVAL; }
    CUDA_TRY(h, cudaSetDevice(h->device));
    if (h->courses_dev) { cudaFree(h->courses_dev); h->courses_dev = nullptr; }
    if (h->dl_dev) { cudaFree(h->dl_dev); h->dl_dev = nullptr; }
    if (h->len_dev) { cudaFree(h->len_dev); h->len_dev = nullptr; }
    const size_t bytes = (size_t)R * 3 * ncmax * sizeof(double);
    CUDA_TRY(h, cudaMalloc(&h->courses_dev, bytes));
    CUDA_TRY(h, cudaMalloc(&h->dl_dev, R * sizeof(double)));
    CUDA_TRY(h, cudaMemcpy(h->courses_dev, courses_host, bytes, cudaMemcpyHostToDevice));
    CUDA_TRY(h, cudaMemcpy(h->dl_dev, dl_host, R * sizeof(double), cudaMemcpyHostToDevice));
    CUDA_TRY(h, cudaMalloc(&h->len_dev, R * sizeof(int)));
    CUDA_TRY(h, cudaMemcpy(h->len_dev, len_host, R * sizeof(int), cudaMemcpyHostToDevice));
    h->len_host.assign(len_host, len_host + R);
    h->goal_host.resize(2 * (size_t)R);
    for (int r = 0; r < R; ++r) {
        const double* c = courses_host + (size_t)r * 3 * ncmax;
        h->goal_host[2 * r] = c[len_host[r] - 1];
        h->goal_host[2 * r + 1] = c[ncmax + len_host[r] - 1];
    }
    // circle centres of every path point (main/lib/trajectories.py:11-55, y offset 0), host libm like numpy
    {
        const mpcb_config& c = h->cfg;
        const double along = c.car_length / 2 - c.car_width / 2, mid = c.wheelbase / 2;   // car_dimensions.py:59-79
        const double offs[2] = {mid + along, mid - along};
        std::vector<double> cc((size_t)R * 4 * ncmax);
        for (int r = 0; r < R; ++r) {
            const double* x = courses_host + (size_t)r * 3 * ncmax, *y = x + ncmax, *th = y + ncmax;
            for (int k = 0; k < 2; ++k)
                for (int i = 0; i < ncmax; ++i) {
                    cc[((size_t)r * 4 + 2 * k) * ncmax + i] = (cos(th[i]) * offs[k] - sin(th[i]) * 0.0) + x[i];
                    cc[((size_t)r * 4 + 2 * k + 1) * ncmax + i] = (sin(th[i]) * offs[k] + cos(th[i]) * 0.0) + y[i];
                }
        }
        if (h->pathcc_dev) { cudaFree(h->pathcc_dev); h->pathcc_dev = nullptr; }
        CUDA_TRY(h, cudaMalloc(&h->pathcc_dev, cc.size() * sizeof(double)));
        CUDA_TRY(h, cudaMemcpy(h->pathcc_dev, cc.data(), cc.size() * sizeof(double), cudaMemcpyHostToDevice));
        // np.cumsum(np.zeros(n) + MAX_ACCEL) (mpc_intersection.py:111-112): sequential float64 sums, index-only
        std::vector<double> ramp((size_t)ncmax);
        double acc = 0.0;
        for (int i = 0; i < ncmax; ++i) { acc = acc + c.max_accel; ramp[i] = acc; }
        if (h->ramp_dev) { cudaFree(h->ramp_dev); h->ramp_dev = nullptr; }
        CUDA_TRY(h, cudaMalloc(&h->ramp_dev, ramp.size() * sizeof(double)));
        CUDA_TRY(h, cudaMemcpy(h->ramp_dev, ramp.data(), ramp.size() * sizeof(double), cudaMemcpyHostToDevice));
    }
    h->R = R; h->ncmax = ncmax;
    return MPCB_OK;
}

int mpcb_step(mpcb_handle* h, int B, const double* state_dev, const int* route_dev, const int* ncourse_dev,
              const double* L_dev, int* target_ind_dev, int* warm_dev, double* oa_dev, double* od_dev,
              double* traj_dev, double* xref_dev, double* di_ai_dev, double* xdev_dev, int* status_dev,
              int* iters_dev, void* stream) {
    if (!h) return MPCB_EINVAL;
    if (B <= 0 || B > h->max_batch || !state_dev || !route_dev || !ncourse_dev || !target_ind_dev || !warm_dev ||
        !oa_dev || !od_dev || !di_ai_dev || !status_dev) { h->err = "mpcb_step: bad arguments"; return MPCB_EINVAL; }
    if (!h->courses_dev) { h->err = "mpcb_step: no courses set"; return MPCB_EINVAL; }
    StepParams p = make_params(h);
    p.B = B; p.state = state_dev; p.route = route_dev; p.ncourse = ncourse_dev; p.L = L_dev;
    p.target_ind = target_ind_dev; p.warm = warm_dev; p.oa = oa_dev; p.od = od_dev; p.traj = traj_dev;
    p.xref = xref_dev; p.di_ai = di_ai_dev; p.xdev = xdev_dev; p.status = status_dev; p.iters = iters_dev;
    cudaStream_t st = (cudaStream_t)stream;
    return dispatch_step(h, p, st);
}

int mpcb_plant_step(mpcb_handle* h, int B, double* state_dev, const double* di_ai_dev, const double* L_dev,
                    const int* active_dev, void* stream) {
    if (!h || B <= 0 || !state_dev || !di_ai_dev) { if (h) h->err = "mpcb_plant_step: bad arguments"; return MPCB_EINVAL; }
    const mpcb_config& c = h->cfg;
    plant_step_kernel<<<(B + 127) / 128, 128, 0, (cudaStream_t)stream>>>(B, state_dev, di_ai_dev, L_dev, active_dev,
                                                                       c.wheelbase, c.dt, c.max_steer, c.max_speed, c.min_speed);
    CUDA_TRY(h, cudaGetLastError());
    h->launches++;
    return MPCB_OK;
}

int mpcb_is_goal(mpcb_handle* h, int B, const double* state_dev, const double* goal_dev, const int* target_ind_dev,
                 const int* ncourse_dev, int* out_dev, void* stream) {
    if (!h || B <= 0 || !state_dev || !goal_dev || !target_ind_dev || !ncourse_dev || !out_dev) { if (h) h->err = "mpcb_is_goal: bad arguments"; return MPCB_EINVAL; }
    is_goal_kernel<<<(B + 127) / 128, 128, 0, (cudaStream_t)stream>>>(B, state_dev, goal_dev, target_ind_dev, ncourse_dev,
                                                                    out_dev, h->cfg.goal_dis, h->cfg.stop_speed);
    CUDA_TRY(h, cudaGetLastError());
    h->launches++;
    return MPCB_OK;
}

// ---- host-buffer variant: one packed H2D copy, one kernel, one packed D2H copy -----------------------
namespace {
struct StageLayout {
    size_t state, L, route, ncourse;            // in only
    size_t oa, od, di_ai, target, warm;         // in/out
    size_t traj, xref, xdev, status, iters;     // out only
    size_t in_end, io_begin, total;
};
inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }
StageLayout stage_layout(int B, int T) {
    StageLayout l;
    size_t o = 0;
    auto put = [&](size_t bytes) { size_t at = o; o = align256(o + bytes); return at; };
    l.state = put(sizeof(double) * 4 * B); l.L = put(sizeof(double) * B);
    l.route = put(sizeof(int) * B); l.ncourse = put(sizeof(int) * B);
    l.io_begin = o;
    l.oa = put(sizeof(double) * B * T); l.od = put(sizeof(double) * B * T); l.di_ai = put(sizeof(double) * 2 * B);
    l.target = put(sizeof(int) * B); l.warm = put(sizeof(int) * B);
    l.in_end = o;
    l.traj = put(sizeof(double) * 4 * (T + 1) * B); l.xref = put(sizeof(double) * 4 * (T + 1) * B);
    l.xdev = put(sizeof(double) * B); l.status = put(sizeof(int) * B); l.iters = put(sizeof(int) * B);
    l.total = o;
    return l;
}
}  // namespace

int mpcb_step_host(mpcb_handle* h, int B, const double* state, const int* route, const int* ncourse,
                   const double* L, int* target_ind, int* warm, double* oa, double* od, double* traj,
                   double* xref, double* di_ai, double* xdev, int* status, int* iters) {
    if (!h) return MPCB_EINVAL;
    if (B <= 0 || B > h->max_batch || !state || !route || !ncourse || !target_ind || !warm || !oa || !od || !di_ai || !status) {
        h->err = "mpcb_step_host: bad arguments"; return MPCB_EINVAL;
    }
    const int T = h->cfg.T;
    if (!h->courses_dev) { h->err = "mpcb_step_host: no courses set"; return MPCB_EINVAL; }
    for (int b = 0; b < B; ++b) {        // a bad index would become an out-of-bounds read inside the kernel
        const int r = route[b];
        if (r < 0 || r >= h->R || ncourse[b] < 1 || ncourse[b] > h->len_host[r] || target_ind[b] < 0) {
            h->err = "mpcb_step_host: problem " + std::to_string(b) + ": route id, ncourse (1..course length) or target_ind (>= 0) out of range";
            return MPCB_EINVAL;
        }
    }
    CUDA_TRY(h, cudaSetDevice(h->device));
    const StageLayout lm = stage_layout(h->max_batch, T);
    if (!h->stage.pinned) {
        CUDA_TRY(h, cudaMallocHost(&h->stage.pinned, lm.total));
        CUDA_TRY(h, cudaMalloc(&h->stage.dev, lm.total));
        h->stage.bytes = lm.total;
    }
    char* hp = (char*)h->stage.pinned;
    char* dp = (char*)h->stage.dev;
    memcpy(hp + lm.state, state, sizeof(double) * 4 * B);
    if (L) memcpy(hp + lm.L, L, sizeof(double) * B);
    memcpy(hp + lm.route, route, sizeof(int) * B);
    memcpy(hp + lm.ncourse, ncourse, sizeof(int) * B);
    memcpy(hp + lm.oa, oa, sizeof(double) * B * T);
    memcpy(hp + lm.od, od, sizeof(double) * B * T);
    memcpy(hp + lm.di_ai, di_ai, sizeof(double) * 2 * B);
    memcpy(hp + lm.target, target_ind, sizeof(int) * B);
    memcpy(hp + lm.warm, warm, sizeof(int) * B);
    cudaStream_t st = h->own_stream;
    CUDA_TRY(h, cudaMemcpyAsync(dp, hp, lm.in_end, cudaMemcpyHostToDevice, st));
    int rc = mpcb_step(h, B, (double*)(dp + lm.state), (int*)(dp + lm.route), (int*)(dp + lm.ncourse),
                       L ? (double*)(dp + lm.L) : nullptr, (int*)(dp + lm.target), (int*)(dp + lm.warm),
                       (double*)(dp + lm.oa), (double*)(dp + lm.od), (double*)(dp + lm.traj), (double*)(dp + lm.xref),
                       (double*)(dp + lm.di_ai), (double*)(dp + lm.xdev), (int*)(dp + lm.status), (int*)(dp + lm.iters), st);
    if (rc != MPCB_OK) return rc;
    CUDA_TRY(h, cudaMemcpyAsync(hp + lm.io_begin, dp + lm.io_begin, lm.total - lm.io_begin, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(h, cudaStreamSynchronize(st));
    memcpy(oa, hp + lm.oa, sizeof(double) * B * T);
    memcpy(od, hp + lm.od, sizeof(double) * B * T);
    memcpy(di_ai, hp + lm.di_ai, sizeof(double) * 2 * B);
    memcpy(target_ind, hp + lm.target, sizeof(int) * B);
    memcpy(warm, hp + lm.warm, sizeof(int) * B);
    if (traj) memcpy(traj, hp + lm.traj, sizeof(double) * 4 * (T + 1) * B);
    if (xref) memcpy(xref, hp + lm.xref, sizeof(double) * 4 * (T + 1) * B);
    if (xdev) memcpy(xdev, hp + lm.xdev, sizeof(double) * B);
    memcpy(status, hp + lm.status, sizeof(int) * B);
    if (iters) memcpy(iters, hp + lm.iters, sizeof(int) * B);
    return MPCB_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------
// resident session: the batch's MPC object state stays on the device between steps
// ---------------------------------------------------------------------------------------------------
namespace {

int dispatch_step(mpcb_handle* h, const StepParams& p, cudaStream_t st) {
    if (h->cfg.jerk_model) return launch_step_long_nt<64, 4, true>(h, p, st);      // main/lib/mpc_jerk.py
    if (p.pw || p.vcut || p.ref_speed != 0.0) {        // controller variants: their own instantiation
        switch (h->cfg.T) {
            case 10: return half_warp_enabled() ? launch_step_half<10, 1, 8, true>(h, p, st) : launch_step<10, 1, 8, true>(h, p, st);
            case 13: return half_warp_enabled() ? launch_step_half<13, 1, 8, true>(h, p, st) : launch_step<13, 1, 8, true>(h, p, st);
            case 20: return launch_step<20, 1, 8, true>(h, p, st);
            default: break;                             // 40, 80: the long-horizon kernel carries them itself
        }
    }
    switch (h->cfg.T) {
        case 10: return half_warp_enabled() ? launch_step_half<10, 1, 8>(h, p, st) : launch_step<10, 1, 8>(h, p, st);
        case 13:
            if (half_warp_enabled()) return launch_step_half<13, 1, 8>(h, p, st);
#ifdef MPCB_TUNE_VARIANTS
            switch (h->variant) {
                case 1: return launch_step<13, 1, 168>(h, p, st);
                case 2: return launch_step<13, 1, 128>(h, p, st);
                default: break;
            }
#endif
            return launch_step<13, 1, 8>(h, p, st);
        case 20:
#ifdef MPCB_TUNE_VARIANTS
            switch (h->variant) {
                case 1: return launch_step<20, 1, 200>(h, p, st);
                case 2: return launch_step<20, 1, 184>(h, p, st);
                case 3: return launch_step<20, 1, 224>(h, p, st);
                case 4: return launch_step<20, 1, 168>(h, p, st);
                default: break;
            }
#endif
            return launch_step<20, 1, 8>(h, p, st);
        case 40:
        case 80: return launch_step_long(h, p, st);
        default: h->err = "mpcb: unsupported horizon"; return MPCB_EINVAL;
    }
}

int session_alloc(mpcb_handle* h) {
    auto& s = h->ses;
    if (s.slab) return MPCB_OK;
    const size_t B = (size_t)h->max_batch, T = (size_t)h->cfg.T;
    size_t o = 0;
    auto put = [&](size_t bytes) { size_t at = o; o = align256(o + bytes); return at; };
    const size_t o_state = put(8 * 4 * B), o_L = put(8 * B), o_oa = put(8 * T * B), o_od = put(8 * T * B),
                 o_traj = put(8 * 4 * (T + 1) * B), o_xref = put(8 * 4 * (T + 1) * B), o_diai = put(8 * 2 * B),
                 o_status = put(4 * B),      // right behind di_ai: the host-buffer step fetches both with ONE copy
                 o_xdev = put(8 * B), o_goal = put(8 * 2 * B), o_route = put(4 * B), o_nc = put(4 * B),
                 o_tgt = put(4 * B), o_warm = put(4 * B), o_iters = put(4 * B),
                 o_done = put(4 * B), o_aidx = put(4 * B), o_coll = put(4 * B), o_cidx = put(4 * B),
                 o_nscr = put(4 * B), o_script = put(8 * 8 * MPCB_CP_MAX_OBS * B), o_obs = put(8 * 6 * MPCB_CP_MAX_OBS * B),
                 o_pw = put(8 * 12 * B), o_lam = put(8 * 8 * T * B), o_vcut = put(4 * B), o_lord = put(4 * B), o_lrank = put(4 * B), o_lhist = put(4 * 65);
    CUDA_TRY(h, cudaMalloc(&s.slab, o));
    CUDA_TRY(h, cudaMemset(s.slab, 0, o));
    char* d = s.slab;
    s.state = (double*)(d + o_state); s.L = (double*)(d + o_L); s.oa = (double*)(d + o_oa); s.od = (double*)(d + o_od);
    s.traj = (double*)(d + o_traj); s.xref = (double*)(d + o_xref); s.di_ai = (double*)(d + o_diai);
    s.xdev = (double*)(d + o_xdev); s.goal = (double*)(d + o_goal); s.route = (int*)(d + o_route);
    s.ncourse = (int*)(d + o_nc); s.target = (int*)(d + o_tgt); s.warm = (int*)(d + o_warm);
    s.status = (int*)(d + o_status); s.iters = (int*)(d + o_iters); s.done = (int*)(d + o_done);
    s.agent_idx = (int*)(d + o_aidx); s.collided = (int*)(d + o_coll); s.coll_idx = (int*)(d + o_cidx);
    s.n_script = (int*)(d + o_nscr); s.script = (double*)(d + o_script); s.obs = (double*)(d + o_obs);
    s.pw = (double*)(d + o_pw); s.vcut = (int*)(d + o_vcut); s.lam_ws = (double*)(d + o_lam);
    s.lpt_order = (int*)(d + o_lord); s.lpt_rank = (int*)(d + o_lrank); s.lpt_hist = (int*)(d + o_lhist);
    o = 0;
    s.pin_state = put(8 * 4 * B); s.pin_diai = put(8 * 2 * B); s.pin_status = put(4 * B);
    s.pin_total = o;
    CUDA_TRY(h, cudaMallocHost(&s.pin, s.pin_total));
    return MPCB_OK;
}

StepParams session_params(mpcb_handle* h) {
    auto& s = h->ses;
    StepParams p = make_params(h);
    p.B = s.B; p.state = s.state; p.route = s.route; p.ncourse = s.ncourse; p.L = s.hasL ? s.L : nullptr;
    p.target_ind = s.target; p.warm = s.warm; p.oa = s.oa; p.od = s.od; p.traj = s.traj; p.xref = s.xref;
    p.di_ai = s.di_ai; p.xdev = s.xdev; p.status = s.status; p.iters = s.iters;
    p.pw = s.has_pw ? s.pw : nullptr; p.vcut = s.has_vcut ? s.vcut : nullptr;
    static const bool ipm_warm = !(getenv("MPCB_IPM_WARM") && atoi(getenv("MPCB_IPM_WARM")) == 0);
    p.lam_ws = ipm_warm ? s.lam_ws : nullptr;
    // longest-first launch order: pays while the launch is MORE than one and no more than a dozen waves of resident warps
    // (8 per SM).  Measured with / without: B = 1024 (under one wave: order is irrelevant, the ordering pass is pure latency)
    // 4.27 / 4.54 M solves/s, 4096: 8.07 / 7.58, 8192: 8.83 / 8.53, 16384: 9.23 / 9.13, 32768: 9.44 / 9.45, 65536: 9.54 / 9.74
    static const int lpt_max_batch = getenv("MPCB_LPT_MAX_BATCH") ? atoi(getenv("MPCB_LPT_MAX_BATCH")) : 16384;
    // The CTA-per-problem kernels (T = 40, 80, mpc_jerk) hold 4 or 1 problems per SM and run for milliseconds: the ordering pass
    // is free there and the tail it removes is not (T = 40, B = 4096: 7 waves of problems that take 5..35 iterations).
    const bool long_kernel = h->cfg.T > 31 || h->cfg.jerk_model;
    const int resident = h->num_sms * (long_kernel ? (h->cfg.T > 40 ? 1 : 4) : 8);
    if (s.B > resident && (long_kernel || s.B <= lpt_max_batch)) { p.lpt_order = s.lpt_order; p.lpt_rank = s.lpt_rank; p.lpt_hist = s.lpt_hist; }
    return p;
}

}  // namespace

extern "C" {

int mpcb_session_begin(mpcb_handle* h, int B, const int* route_host, const int* ncourse_host, const double* L_host,
                       const int* target_ind_host) {
    if (!h || B <= 0 || B > h->max_batch || !route_host) { if (h) h->err = "mpcb_session_begin: bad arguments"; return MPCB_EINVAL; }
    if (!h->courses_dev) { h->err = "mpcb_session_begin: no courses set"; return MPCB_EINVAL; }
    CUDA_TRY(h, cudaSetDevice(h->device));
    if (h->last_stream_set) CUDA_TRY(h, cudaStreamSynchronize(h->last_stream));
    int rc = session_alloc(h);
    if (rc != MPCB_OK) return rc;
    auto& s = h->ses;
    std::vector<int> nc(B);
    std::vector<double> goal(2 * (size_t)B);
    for (int b = 0; b < B; ++b) {
        const int r = route_host[b];
        if (r < 0 || r >= h->R) { h->err = "mpcb_session_begin: route id out of range"; return MPCB_EINVAL; }
        nc[b] = ncourse_host ? ncourse_host[b] : h->len_host[r];
        if (nc[b] < 1 || nc[b] > h->len_host[r]) { h->err = "mpcb_session_begin: ncourse out of range"; return MPCB_EINVAL; }
        goal[2 * b] = h->goal_host[2 * r]; goal[2 * b + 1] = h->goal_host[2 * r + 1];
    }
    if (target_ind_host)
        for (int b = 0; b < B; ++b)
            if (target_ind_host[b] < 0) { h->err = "mpcb_session_begin: target_ind must be >= 0"; return MPCB_EINVAL; }
    s.B = B; s.hasL = L_host != nullptr; s.has_script = false; s.has_pw = false; s.has_vcut = false;
    s.route_host.assign(route_host, route_host + B);
    cudaStream_t st = h->own_stream;
    CUDA_TRY(h, cudaMemcpyAsync(s.route, route_host, 4 * (size_t)B, cudaMemcpyHostToDevice, st));
    CUDA_TRY(h, cudaMemcpyAsync(s.ncourse, nc.data(), 4 * (size_t)B, cudaMemcpyHostToDevice, st));
    CUDA_TRY(h, cudaMemcpyAsync(s.goal, goal.data(), 16 * (size_t)B, cudaMemcpyHostToDevice, st));
    if (L_host) CUDA_TRY(h, cudaMemcpyAsync(s.L, L_host, 8 * (size_t)B, cudaMemcpyHostToDevice, st));
    if (target_ind_host) CUDA_TRY(h, cudaMemcpyAsync(s.target, target_ind_host, 4 * (size_t)B, cudaMemcpyHostToDevice, st));
    else CUDA_TRY(h, cudaMemsetAsync(s.target, 0, 4 * (size_t)B, st));
    CUDA_TRY(h, cudaMemsetAsync(s.warm, 0, 4 * (size_t)B, st));         // oa = odelta = None (mpc.py:265-266)
    {   // launch order of the first step: slot i solves problem i
        std::vector<int> iota((size_t)B);
        for (int b = 0; b < B; ++b) iota[b] = b;
        CUDA_TRY(h, cudaMemcpyAsync(s.lpt_order, iota.data(), 4 * (size_t)B, cudaMemcpyHostToDevice, st));
        CUDA_TRY(h, cudaMemsetAsync(s.lpt_hist, 0, 4 * 65, st));
        CUDA_TRY(h, cudaStreamSynchronize(st));
    }
    CUDA_TRY(h, cudaMemsetAsync(s.done, 0, 4 * (size_t)B, st));
    CUDA_TRY(h, cudaMemsetAsync(s.agent_idx, 0, 4 * (size_t)B, st));       // traj_agent_idx = 0 (mpc_intersection.py:78)
    CUDA_TRY(h, cudaMemsetAsync(s.collided, 0, 4 * (size_t)B, st));
    CUDA_TRY(h, cudaMemsetAsync(s.coll_idx, 0xff, 4 * (size_t)B, st));
    CUDA_TRY(h, cudaMemsetAsync(s.di_ai, 0, 16 * (size_t)B, st));       // di = ai = 0.0 (mpc.py:270-271)
    CUDA_TRY(h, cudaMemsetAsync(s.oa, 0, 8 * (size_t)B * h->cfg.T, st));
    CUDA_TRY(h, cudaMemsetAsync(s.od, 0, 8 * (size_t)B * h->cfg.T, st));
    CUDA_TRY(h, cudaStreamSynchronize(st));
    return MPCB_OK;
}

int mpcb_session_set_ncourse_host(mpcb_handle* h, const int* ncourse_host) {
    if (!h || !ncourse_host || h->ses.B <= 0) { if (h) h->err = "mpcb_session_set_ncourse_host: no session"; return MPCB_EINVAL; }
    for (int b = 0; b < h->ses.B; ++b)
        if (ncourse_host[b] < 1 || ncourse_host[b] > h->len_host[h->ses.route_host[b]]) {
            h->err = "mpcb_session_set_ncourse_host: problem " + std::to_string(b) + ": ncourse outside 1..course length"; return MPCB_EINVAL;
        }
    CUDA_TRY(h, cudaSetDevice(h->device));
    if (h->last_stream_set) CUDA_TRY(h, cudaStreamSynchronize(h->last_stream));
    CUDA_TRY(h, cudaMemcpyAsync(h->ses.ncourse, ncourse_host, 4 * (size_t)h->ses.B, cudaMemcpyHostToDevice, h->own_stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->own_stream));
    return MPCB_OK;
}

int mpcb_session_set_weights_host(mpcb_handle* h, const double* weights_host) {
    if (!h || h->ses.B <= 0) { if (h) h->err = "mpcb_session_set_weights_host: no session"; return MPCB_EINVAL; }
    auto& s = h->ses;
    CUDA_TRY(h, cudaSetDevice(h->device));
    if (h->last_stream_set) CUDA_TRY(h, cudaStreamSynchronize(h->last_stream));
    if (!weights_host) { s.has_pw = false; return MPCB_OK; }
    std::vector<double> w(weights_host, weights_host + 12 * (size_t)s.B);
    for (size_t b = 0; b < (size_t)s.B; ++b)
        for (int i = 8; i < 12; ++i) w[12 * b + i] *= h->cfg.T;          // Qf = np.diag(config['Qf']) * T (mpc.py:25)
    CUDA_TRY(h, cudaMemcpyAsync(s.pw, w.data(), 8 * w.size(), cudaMemcpyHostToDevice, h->own_stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->own_stream));
    s.has_pw = true;
    return MPCB_OK;
}

int mpcb_session_set_speed_cutoff_host(mpcb_handle* h, const int* cutoff_idx_host) {
    if (!h || h->ses.B <= 0) { if (h) h->err = "mpcb_session_set_speed_cutoff_host: no session"; return MPCB_EINVAL; }
    auto& s = h->ses;
    CUDA_TRY(h, cudaSetDevice(h->device));
    if (h->last_stream_set) CUDA_TRY(h, cudaStreamSynchronize(h->last_stream));
    if (!cutoff_idx_host) { s.has_vcut = false; return MPCB_OK; }
    // 999 is the reference's "no cut" sentinel (mpc_with_speed.py:276,281), not an index: a course longer than 999 points keeps
    // its speed reference past point 999
    std::vector<int> cut(cutoff_idx_host, cutoff_idx_host + s.B);
    for (int& c : cut) if (c == 999 || c < 0) c = 0x7fffffff;
    CUDA_TRY(h, cudaMemcpyAsync(s.vcut, cut.data(), 4 * (size_t)s.B, cudaMemcpyHostToDevice, h->own_stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->own_stream));
    s.has_vcut = true;
    return MPCB_OK;
}

int mpcb_session_step_host(mpcb_handle* h, const double* state_host, double* di_ai_host, int* status_host) {
    if (!h || !state_host || !di_ai_host || h->ses.B <= 0) { if (h) h->err = "mpcb_session_step_host: bad arguments / no session"; return MPCB_EINVAL; }
    auto& s = h->ses;
    const size_t B = (size_t)s.B;
    CUDA_TRY(h, cudaSetDevice(h->device));
    if (h->last_stream_set) CUDA_TRY(h, cudaStreamSynchronize(h->last_stream));
    cudaStream_t st = h->own_stream;
    memcpy(s.pin + s.pin_state, state_host, 32 * B);
    CUDA_TRY(h, cudaMemcpyAsync(s.state, s.pin + s.pin_state, 32 * B, cudaMemcpyHostToDevice, st));
    int rc = dispatch_step(h, session_params(h), st);
    if (rc != MPCB_OK) return rc;
    if ((char*)s.status - (char*)s.di_ai == (ptrdiff_t)(s.pin_status - s.pin_diai) && s.pin_status - s.pin_diai < 16 * B + 256) {
        // full batch: di_ai and status are neighbours in the slab and in the pinned buffer -- one copy, one DMA latency
        CUDA_TRY(h, cudaMemcpyAsync(s.pin + s.pin_diai, s.di_ai, (s.pin_status - s.pin_diai) + 4 * B, cudaMemcpyDeviceToHost, st));
    } else {
        CUDA_TRY(h, cudaMemcpyAsync(s.pin + s.pin_diai, s.di_ai, 16 * B, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(h, cudaMemcpyAsync(s.pin + s.pin_status, s.status, 4 * B, cudaMemcpyDeviceToHost, st));
    }
    CUDA_TRY(h, cudaStreamSynchronize(st));
    memcpy(di_ai_host, s.pin + s.pin_diai, 16 * B);
    if (status_host) memcpy(status_host, s.pin + s.pin_status, 4 * B);
    return MPCB_OK;
}

// One MPC.step of the resident session with everything a caller of the reference reads afterwards (mpc.py:289-292 and
// the attribute reads of main/scenarios/mpc_intersection.py:297-302) in ONE synchronisation: inputs staged by asynchronous
// copies, the kernel, then one asynchronous D2H per requested field into pinned memory and a single stream wait.
int mpcb_session_step_all_host(mpcb_handle* h, const double* state_host, const int* target_ind_host, const int* warm_host,
                               double* di_ai_host, int* status_host, int* target_ind_out, int* iters_out, double* oa_out,
                               double* od_out, double* traj_out, double* xref_out) {
    if (!h || !state_host || !di_ai_host || h->ses.B <= 0) { if (h) h->err = "mpcb_session_step_all_host: bad arguments / no session"; return MPCB_EINVAL; }
    auto& s = h->ses;
    const size_t B = (size_t)s.B, T = (size_t)h->cfg.T;
    if (target_ind_host)
        for (size_t b = 0; b < B; ++b)
            if (target_ind_host[b] < 0) { h->err = "mpcb_session_step_all_host: target_ind must be >= 0"; return MPCB_EINVAL; }
    CUDA_TRY(h, cudaSetDevice(h->device));
    if (h->last_stream_set) CUDA_TRY(h, cudaStreamSynchronize(h->last_stream));
    size_t o = 0;
    auto put = [&](size_t bytes) { size_t at = o; o = align256(o + bytes); return at; };
    const size_t p_state = put(32 * B), p_tgt = put(4 * B), p_warm = put(4 * B), p_diai = put(16 * B), p_status = put(4 * B),
                 p_iters = put(4 * B), p_oa = put(8 * T * B), p_od = put(8 * T * B), p_traj = put(32 * (T + 1) * B), p_xref = put(32 * (T + 1) * B);
    if (s.pin_all_bytes < o) {
        if (s.pin_all) cudaFreeHost(s.pin_all);
        s.pin_all = nullptr; s.pin_all_bytes = 0;
        CUDA_TRY(h, cudaMallocHost(&s.pin_all, o));
        s.pin_all_bytes = o;
    }
    char* pin = s.pin_all;
    cudaStream_t st = h->own_stream;
    memcpy(pin + p_state, state_host, 32 * B);
    CUDA_TRY(h, cudaMemcpyAsync(s.state, pin + p_state, 32 * B, cudaMemcpyHostToDevice, st));
    if (target_ind_host) { memcpy(pin + p_tgt, target_ind_host, 4 * B); CUDA_TRY(h, cudaMemcpyAsync(s.target, pin + p_tgt, 4 * B, cudaMemcpyHostToDevice, st)); }
    if (warm_host) { memcpy(pin + p_warm, warm_host, 4 * B); CUDA_TRY(h, cudaMemcpyAsync(s.warm, pin + p_warm, 4 * B, cudaMemcpyHostToDevice, st)); }
    int rc = dispatch_step(h, session_params(h), st);
    if (rc != MPCB_OK) return rc;
    auto down = [&](void* want, size_t at, const void* dev, size_t bytes) {
        return want ? cudaMemcpyAsync(pin + at, dev, bytes, cudaMemcpyDeviceToHost, st) : cudaSuccess;
    };
    CUDA_TRY(h, down(di_ai_host, p_diai, s.di_ai, 16 * B));
    CUDA_TRY(h, down(status_host, p_status, s.status, 4 * B));
    CUDA_TRY(h, down(target_ind_out, p_tgt, s.target, 4 * B));
    CUDA_TRY(h, down(iters_out, p_iters, s.iters, 4 * B));
    CUDA_TRY(h, down(oa_out, p_oa, s.oa, 8 * T * B));
    CUDA_TRY(h, down(od_out, p_od, s.od, 8 * T * B));
    CUDA_TRY(h, down(traj_out, p_traj, s.traj, 32 * (T + 1) * B));
    CUDA_TRY(h, down(xref_out, p_xref, s.xref, 32 * (T + 1) * B));
    CUDA_TRY(h, cudaStreamSynchronize(st));
    auto out = [&](void* want, size_t at, size_t bytes) { if (want) memcpy(want, pin + at, bytes); };
    out(di_ai_host, p_diai, 16 * B); out(status_host, p_status, 4 * B); out(target_ind_out, p_tgt, 4 * B); out(iters_out, p_iters, 4 * B);
    out(oa_out, p_oa, 8 * T * B); out(od_out, p_od, 8 * T * B); out(traj_out, p_traj, 32 * (T + 1) * B); out(xref_out, p_xref, 32 * (T + 1) * B);
    return MPCB_OK;
}

int mpcb_session_step_dev(mpcb_handle* h, void* stream) {
    if (!h || h->ses.B <= 0) { if (h) h->err = "mpcb_session_step_dev: no session"; return MPCB_EINVAL; }
    h->last_stream = (cudaStream_t)stream; h->last_stream_set = true;
    return dispatch_step(h, session_params(h), (cudaStream_t)stream);
}

int mpcb_session_closed_loop_dev(mpcb_handle* h, int nsteps, double* history_dev, int flags, void* stream) {
    if (!h || h->ses.B <= 0 || nsteps <= 0) { if (h) h->err = "mpcb_session_closed_loop_dev: bad arguments / no session"; return MPCB_EINVAL; }
    auto& s = h->ses;
    h->last_stream = (cudaStream_t)stream; h->last_stream_set = true;
    StepParams p = session_params(h);
    p.done = s.done; p.goal = s.goal; p.state_out = s.state; p.fuse_goal = (flags & MPCB_CL_NO_GOAL_TEST) ? 0 : 1; p.fuse_plant = 1;
    for (int k = 0; k < nsteps; ++k) {
        p.history = history_dev ? history_dev + (size_t)k * 8 * s.B : nullptr;
        int rc = dispatch_step(h, p, (cudaStream_t)stream);
        if (rc != MPCB_OK) return rc;
    }
    return MPCB_OK;
}

void* mpcb_session_device_ptr(mpcb_handle* h, int what) {
    if (!h || !h->ses.slab) return nullptr;
    auto& s = h->ses;
    switch (what) {
        case MPCB_F_STATE: return s.state;   case MPCB_F_OA: return s.oa;         case MPCB_F_OD: return s.od;
        case MPCB_F_TRAJ: return s.traj;     case MPCB_F_XREF: return s.xref;     case MPCB_F_DI_AI: return s.di_ai;
        case MPCB_F_XDEV: return s.xdev;     case MPCB_F_TARGET: return s.target; case MPCB_F_WARM: return s.warm;
        case MPCB_F_STATUS: return s.status; case MPCB_F_ITERS: return s.iters;   case MPCB_F_DONE: return s.done;
        case MPCB_F_NCOURSE: return s.ncourse; case MPCB_F_ROUTE: return s.route; case MPCB_F_GOAL: return s.goal;
        case MPCB_F_AGENT_IDX: return s.agent_idx; case MPCB_F_COLLIDED: return s.collided; case MPCB_F_COLL_IDX: return s.coll_idx;
        case MPCB_F_SCRIPT: return s.script; case MPCB_F_OBS: return s.obs; case MPCB_F_NOBS: return s.n_script;
        case MPCB_F_LAUNCH_ORDER: return s.lpt_order;
        default: return nullptr;
    }
}

static size_t field_bytes(const mpcb_handle* h, int what) {
    const size_t B = (size_t)h->ses.B, T = (size_t)h->cfg.T;
    switch (what) {
        case MPCB_F_STATE: return 32 * B;    case MPCB_F_OA: case MPCB_F_OD: return 8 * T * B;
        case MPCB_F_TRAJ: case MPCB_F_XREF: return 32 * (T + 1) * B;
        case MPCB_F_DI_AI: case MPCB_F_GOAL: return 16 * B;  case MPCB_F_XDEV: return 8 * B;
        case MPCB_F_TARGET: case MPCB_F_WARM: case MPCB_F_STATUS: case MPCB_F_ITERS: case MPCB_F_DONE:
        case MPCB_F_NCOURSE: case MPCB_F_ROUTE: case MPCB_F_AGENT_IDX: case MPCB_F_COLLIDED: case MPCB_F_COLL_IDX:
        case MPCB_F_NOBS: case MPCB_F_LAUNCH_ORDER: return 4 * B;
        case MPCB_F_SCRIPT: return 8 * 8 * MPCB_CP_MAX_OBS * B;
        case MPCB_F_OBS: return 8 * 6 * MPCB_CP_MAX_OBS * B;
        default: return 0;
    }
}

long long mpcb_session_field_bytes(const mpcb_handle* h, int what) {
    if (!h || h->ses.B <= 0) return MPCB_EINVAL;
    return (long long)field_bytes(h, what);
}

int mpcb_session_fetch_host(mpcb_handle* h, int what, void* dst, long long bytes) {
    if (!h || !dst || h->ses.B <= 0) { if (h) h->err = "mpcb_session_fetch_host: bad arguments / no session"; return MPCB_EINVAL; }
    void* src = mpcb_session_device_ptr(h, what);
    const size_t need = field_bytes(h, what);
    if (!src || need == 0 || (size_t)bytes != need) { h->err = "mpcb_session_fetch_host: unknown field or wrong size"; return MPCB_EINVAL; }
    CUDA_TRY(h, cudaSetDevice(h->device));
    if (h->last_stream_set) CUDA_TRY(h, cudaStreamSynchronize(h->last_stream));   // order after asynchronous session work
    CUDA_TRY(h, cudaMemcpyAsync(dst, src, need, cudaMemcpyDeviceToHost, h->own_stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->own_stream));
    return MPCB_OK;
}

int mpcb_session_store_host(mpcb_handle* h, int what, const void* src, long long bytes) {
    if (!h || !src || h->ses.B <= 0) { if (h) h->err = "mpcb_session_store_host: bad arguments / no session"; return MPCB_EINVAL; }
    void* dst = mpcb_session_device_ptr(h, what);
    const size_t need = field_bytes(h, what);
    if (!dst || need == 0 || (size_t)bytes != need) { h->err = "mpcb_session_store_host: unknown field or wrong size"; return MPCB_EINVAL; }
    if (what == MPCB_F_ROUTE || what == MPCB_F_NCOURSE || what == MPCB_F_TARGET) {      // indices the kernels trust
        const int* v = (const int*)src;
        auto& rh = h->ses.route_host;
        for (int b = 0; b < h->ses.B; ++b) {
            const bool ok = what == MPCB_F_ROUTE ? (v[b] >= 0 && v[b] < h->R)
                          : what == MPCB_F_NCOURSE ? (v[b] >= 1 && v[b] <= h->len_host[rh[b]]) : v[b] >= 0;
            if (!ok) { h->err = "mpcb_session_store_host: problem " + std::to_string(b) + ": index out of range"; return MPCB_EINVAL; }
        }
        if (what == MPCB_F_ROUTE) rh.assign(v, v + h->ses.B);
    }
    CUDA_TRY(h, cudaSetDevice(h->device));
    if (h->last_stream_set) CUDA_TRY(h, cudaStreamSynchronize(h->last_stream));   // order after asynchronous session work
    CUDA_TRY(h, cudaMemcpyAsync(dst, src, need, cudaMemcpyHostToDevice, h->own_stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->own_stream));
    return MPCB_OK;
}

}  // extern "C"

extern "C" {

}  // extern "C"

// ---------------------------------------------------------------------------------------------------
// row A9: inter-vehicle coupling
// ---------------------------------------------------------------------------------------------------
namespace {
int launch_coupling(mpcb_handle* h, CouplingParams& p, cudaStream_t st) {
    const mpcb_config& c = h->cfg;
    p.ncmax = h->ncmax; p.courses = h->courses_dev; p.pathcc = h->pathcc_dev; p.accel_ramp = h->ramp_dev; p.course_len = h->len_dev; p.dl = h->dl_dev;
    const double along = c.car_length / 2 - c.car_width / 2, mid = c.wheelbase / 2;
    p.circ[0] = mid + along; p.circ[1] = mid - along;
    p.radius = c.car_width / sqrt(2.0);                          // width / 2**.5 (car_dimensions.py:64-65)
    p.wheelbase = c.wheelbase; p.dt = c.dt; p.max_accel = c.max_accel; p.max_speed = c.max_speed;
    p.frame_window = c.frame_window;
    p.n_frames = (int)ceil(c.coupling_horizon / c.dt);           // len(np.arange(0, horizon, dt))
    if (p.n_frames < 1 || p.n_frames > MPCB_CP_FRAMES || p.frame_window < 0 || p.frame_window > 64) {
        h->err = "mpcb_coupling: coupling_horizon/dt must give 1..35 frames and frame_window 0..64"; return MPCB_EINVAL;
    }
    constexpr int WARPS = 4;
    const size_t smem = sizeof(mpcb::CouplingSmem) * WARPS;
    auto kern = mpcb::coupling_kernel<WARPS>;
    static thread_local int configured_dev = -1;
    if (configured_dev != h->device) {
        CUDA_TRY(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured_dev = h->device;
    }
    kern<<<(p.B + WARPS - 1) / WARPS, WARPS * 32, smem, st>>>(p);
    CUDA_TRY(h, cudaGetLastError());
    h->launches++;
    return MPCB_OK;
}
}  // namespace

extern "C" {

int mpcb_coupling(mpcb_handle* h, int B, const double* state_dev, const int* route_dev, const double* obs_dev,
                  const int* n_obs_dev, int* done_dev, int* agent_idx_dev, int* ncourse_dev, int* collided_dev,
                  int* coll_idx_dev, double* coll_xy_dev, int* n_res_dev, int* flags_dev, int vehicles_per_scenario,
                  void* stream) {
    if (!h) return MPCB_EINVAL;
    if (B <= 0 || !state_dev || !route_dev || !agent_idx_dev || !ncourse_dev) { h->err = "mpcb_coupling: bad arguments"; return MPCB_EINVAL; }
    if (!h->courses_dev) { h->err = "mpcb_coupling: no courses set"; return MPCB_EINVAL; }
    if (vehicles_per_scenario < 0 || vehicles_per_scenario > MPCB_CP_MAX_OBS + 1 ||
        (vehicles_per_scenario > 0 && B % vehicles_per_scenario != 0)) {
        h->err = "mpcb_coupling: vehicles_per_scenario must be 0..9 and divide B"; return MPCB_EINVAL;
    }
    if (vehicles_per_scenario == 0 && n_obs_dev && !obs_dev) { h->err = "mpcb_coupling: n_obs without obs"; return MPCB_EINVAL; }
    CouplingParams p;
    memset(&p, 0, sizeof(p));
    p.B = B; p.vehicles_per_scenario = vehicles_per_scenario; p.state = state_dev; p.route = route_dev; p.obs = obs_dev;
    p.n_obs = n_obs_dev; p.done = done_dev; p.agent_idx = agent_idx_dev; p.ncourse = ncourse_dev; p.collided = collided_dev;
    p.coll_idx = coll_idx_dev; p.coll_xy = coll_xy_dev; p.n_res = n_res_dev; p.flags = flags_dev;
    return launch_coupling(h, p, (cudaStream_t)stream);
}

int mpcb_coupling_host(mpcb_handle* h, int B, const double* state, const int* route, const double* obs,
                       const int* n_obs, int* agent_idx, int* ncourse, int* collided, int* coll_idx, double* coll_xy,
                       int* n_res, int* flags, int vehicles_per_scenario) {
    if (!h) return MPCB_EINVAL;
    if (B <= 0 || !state || !route || !agent_idx || !ncourse) { h->err = "mpcb_coupling_host: bad arguments"; return MPCB_EINVAL; }
    CUDA_TRY(h, cudaSetDevice(h->device));
    const size_t nB = (size_t)B;
    size_t o = 0;
    auto put = [&](size_t bytes) { size_t at = o; o = align256(o + bytes); return at; };
    const size_t o_state = put(32 * nB), o_route = put(4 * nB), o_obs = put(8 * 6 * MPCB_CP_MAX_OBS * nB), o_nobs = put(4 * nB),
                 o_aidx = put(4 * nB), o_nc = put(4 * nB), o_coll = put(4 * nB), o_cidx = put(4 * nB), o_cxy = put(16 * nB),
                 o_nres = put(4 * nB), o_flags = put(4 * nB);
    char* d = nullptr;
    CUDA_TRY(h, cudaMalloc(&d, o));
    cudaStream_t st = h->own_stream;
    int rc = MPCB_OK;
    auto up = [&](size_t off, const void* src, size_t bytes) { if (src && cudaMemcpyAsync(d + off, src, bytes, cudaMemcpyHostToDevice, st) != cudaSuccess) rc = MPCB_ECUDA; };
    up(o_state, state, 32 * nB); up(o_route, route, 4 * nB); up(o_obs, obs, 8 * 6 * MPCB_CP_MAX_OBS * nB);
    up(o_nobs, n_obs, 4 * nB); up(o_aidx, agent_idx, 4 * nB); up(o_nc, ncourse, 4 * nB);
    cudaMemsetAsync(d + o_flags, 0, 4 * nB, st);
    if (rc == MPCB_OK)
        rc = mpcb_coupling(h, B, (double*)(d + o_state), (int*)(d + o_route), obs ? (double*)(d + o_obs) : nullptr,
                           n_obs ? (int*)(d + o_nobs) : nullptr, nullptr, (int*)(d + o_aidx), (int*)(d + o_nc),
                           (int*)(d + o_coll), (int*)(d + o_cidx), (double*)(d + o_cxy), (int*)(d + o_nres),
                           (int*)(d + o_flags), vehicles_per_scenario, st);
    auto down = [&](void* dst, size_t off, size_t bytes) { if (dst && rc == MPCB_OK && cudaMemcpyAsync(dst, d + off, bytes, cudaMemcpyDeviceToHost, st) != cudaSuccess) rc = MPCB_ECUDA; };
    down(agent_idx, o_aidx, 4 * nB); down(ncourse, o_nc, 4 * nB); down(collided, o_coll, 4 * nB); down(coll_idx, o_cidx, 4 * nB);
    down(coll_xy, o_cxy, 16 * nB); down(n_res, o_nres, 4 * nB); down(flags, o_flags, 4 * nB);
    if (cudaStreamSynchronize(st) != cudaSuccess && rc == MPCB_OK) { h->err = "mpcb_coupling_host: CUDA failure"; rc = MPCB_ECUDA; }
    cudaFree(d);
    return rc;
}

// Coupled closed loop shared by the multi-agent sweep (obstacles = the other egos of a scenario) and the scripted
// cross-traffic sweep (obstacles = MovingObstacleTIntersection scripts stepped inside the coupling kernel).
static int coupled_loop(mpcb_handle* h, int nsteps, int vehicles_per_scenario, bool scripted, double* history_dev,
                        double* obs_history_dev, int flags, cudaStream_t stream) {
    auto& s = h->ses;
    h->last_stream = stream; h->last_stream_set = true;
    StepParams p = session_params(h);
    p.done = s.done; p.goal = s.goal; p.state_out = s.state; p.fuse_goal = 0; p.fuse_plant = 1;
    for (int k = 0; k < nsteps; ++k) {
        // loop head of the reference (mpc_intersection.py:97): is_goal against the path as cut by the previous step,
        // evaluated inside the coupling kernel; then the cut; then mpc.step + simulation.step.  All egos read
        // start-of-step states (Jacobi order); finished egos stay where they are and still act as obstacles.
        CouplingParams cp;
        memset(&cp, 0, sizeof(cp));
        cp.B = s.B; cp.vehicles_per_scenario = vehicles_per_scenario; cp.state = s.state; cp.route = s.route;
        cp.done = s.done; cp.agent_idx = s.agent_idx; cp.ncourse = s.ncourse; cp.collided = s.collided; cp.coll_idx = s.coll_idx;
        cp.goal = s.goal; cp.target_ind = s.target; cp.fuse_goal = (flags & MPCB_CL_NO_GOAL_TEST) ? 0 : 1;
        cp.goal_dis = h->cfg.goal_dis; cp.stop_speed = h->cfg.stop_speed;
        if (scripted) {
            cp.script = s.script; cp.n_obs = s.n_script;
            cp.obs_out = obs_history_dev ? obs_history_dev + (size_t)k * 6 * MPCB_CP_MAX_OBS * s.B : s.obs;
        }
        int rc = launch_coupling(h, cp, stream);
        if (rc != MPCB_OK) return rc;
        p.history = history_dev ? history_dev + (size_t)k * 8 * s.B : nullptr;
        rc = dispatch_step(h, p, stream);
        if (rc != MPCB_OK) return rc;
    }
    return MPCB_OK;
}

int mpcb_session_multi_agent_dev(mpcb_handle* h, int nsteps, int vehicles_per_scenario, double* history_dev, int flags,
                                 void* stream) {
    if (!h || h->ses.B <= 0 || nsteps <= 0 || vehicles_per_scenario < 1) { if (h) h->err = "mpcb_session_multi_agent_dev: bad arguments / no session"; return MPCB_EINVAL; }
    if (vehicles_per_scenario > MPCB_CP_MAX_OBS + 1 || h->ses.B % vehicles_per_scenario != 0) {
        h->err = "mpcb_session_multi_agent_dev: vehicles_per_scenario must be 1..9 and divide B"; return MPCB_EINVAL;
    }
    return coupled_loop(h, nsteps, vehicles_per_scenario, false, history_dev, nullptr, flags, (cudaStream_t)stream);
}

int mpcb_session_set_scripted_host(mpcb_handle* h, const double* script_host, const int* n_obs_host) {
    if (!h || h->ses.B <= 0 || !script_host || !n_obs_host) { if (h) h->err = "mpcb_session_set_scripted_host: bad arguments / no session"; return MPCB_EINVAL; }
    auto& s = h->ses;
    const size_t B = (size_t)s.B;
    std::vector<double> st(B * MPCB_CP_MAX_OBS * 8, 0.0);
    for (size_t b = 0; b < B; ++b) {
        if (n_obs_host[b] < 0 || n_obs_host[b] > MPCB_CP_MAX_OBS) { h->err = "mpcb_session_set_scripted_host: n_obs out of range"; return MPCB_EINVAL; }
        for (int o = 0; o < n_obs_host[b]; ++o) {
            const double* in = script_host + (b * MPCB_CP_MAX_OBS + o) * 4;      // direction, turning, speed, offset
            double* out = &st[(b * MPCB_CP_MAX_OBS + o) * 8];
            const bool right = in[0] >= 0.0;                                       // moving_obstacles.py:177
            out[0] = right ? -30.0 : 30.0;                                         // :186-196
            out[1] = right ? -3.0 : 3.0;
            out[2] = right ? 0.0 : 3.141592653589793;
            out[3] = 0.0;                                                          // counter (:185)
            out[4] = right ? 1.0 : -1.0;
            out[5] = in[1] != 0.0 ? 1.0 : 0.0;
            out[6] = in[2];
            out[7] = (in[3] > 0.0) ? in[3] : 0.0;                                  // offset None unless > 0 (:181)
        }
    }
    CUDA_TRY(h, cudaSetDevice(h->device));
    if (h->last_stream_set) CUDA_TRY(h, cudaStreamSynchronize(h->last_stream));
    CUDA_TRY(h, cudaMemcpyAsync(s.script, st.data(), 8 * st.size(), cudaMemcpyHostToDevice, h->own_stream));
    CUDA_TRY(h, cudaMemcpyAsync(s.n_script, n_obs_host, 4 * B, cudaMemcpyHostToDevice, h->own_stream));
    CUDA_TRY(h, cudaMemsetAsync(s.obs, 0, 8 * 6 * MPCB_CP_MAX_OBS * B, h->own_stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->own_stream));
    s.has_script = true;
    return MPCB_OK;
}

int mpcb_session_scripted_loop_dev(mpcb_handle* h, int nsteps, double* history_dev, double* obs_history_dev, int flags,
                                   void* stream) {
    if (!h || h->ses.B <= 0 || nsteps <= 0) { if (h) h->err = "mpcb_session_scripted_loop_dev: bad arguments / no session"; return MPCB_EINVAL; }
    if (!h->ses.has_script) { h->err = "mpcb_session_scripted_loop_dev: call mpcb_session_set_scripted_host first"; return MPCB_EINVAL; }
    return coupled_loop(h, nsteps, 0, true, history_dev, obs_history_dev, flags, (cudaStream_t)stream);
}

}  // extern "C"

extern "C" {

long long mpcb_launch_count(const mpcb_handle* h) { return h ? h->launches : 0; }

int mpcb_measure_fp64_peak(int device, double* tflops_out) {
    if (!tflops_out) return MPCB_EINVAL;
    if (cudaSetDevice(device) != cudaSuccess) return MPCB_ENODEV;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return MPCB_ECUDA;
    const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 1 << 15;
    double* out = nullptr;
    if (cudaMalloc(&out, sizeof(double) * blocks * threads) != cudaSuccess) return MPCB_ENOMEM;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    double best = 0.0;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        fp64_peak_kernel<<<blocks, threads>>>(out, iters, 1.0 + rep);
        cudaEventRecord(e1);
        if (cudaEventSynchronize(e1) != cudaSuccess) { cudaFree(out); return MPCB_ECUDA; }
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        const double flops = 2.0 * 8.0 * (double)iters * blocks * threads;
        if (rep > 0) best = fmax(best, flops / (ms * 1e-3) / 1e12);
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(out);
    *tflops_out = best;
    return MPCB_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------
// allocation-free step: mpcb_step with the longest-first launch order, its scratch in static device memory
// ---------------------------------------------------------------------------------------------------
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops).  The step kernel's launch-order scratch (histogram + finished-warp
// counter, per-problem ranks, slot -> problem table; see lpt_finish in mpcb_step.cuh) normally lives in the session slab
// that mpcb_session_begin allocates.  Here it is a file-scope array that every GPU loading the library holds until exit:
// words [0, 65) histogram + counter, word 128 the batch size of the last call, [256, 256 + MAX) order, then MAX ranks.
namespace {
constexpr int kStaticScratchHead = 256;
constexpr size_t kStaticScratchWords = kStaticScratchHead + 2 * (size_t)MPCB_STATIC_MAX_BATCH;
static __device__ __align__(256) int g_step_scratch[kStaticScratchWords];      // 129 KB per device

// One CTA: the histogram and the counter must start at zero on every call; the slot -> problem table of the last call is
// kept (that is the point: problems keep their cost from one control step to the next) unless the batch size changed,
// in which case it is no permutation of this batch and is reset to the identity.
__global__ void static_scratch_prepare(int* __restrict__ scr, int B) {
    const bool fresh = scr[128] != B;
    if (threadIdx.x < 65) scr[threadIdx.x] = 0;
    if (fresh)
        for (int i = threadIdx.x; i < B; i += blockDim.x) scr[kStaticScratchHead + i] = i;
    __syncthreads();
    if (threadIdx.x == 0) scr[128] = B;
}

constexpr int kMaxDevices = 64;
int* g_step_scratch_ptr[kMaxDevices] = {};      // device address of g_step_scratch on each device (host code cannot use the name)

// once per device, with that device current: the array's name is a host-side stand-in, only this gives the address
int static_scratch_ptr(mpcb_handle* h, int** out) {
    if (h->device < 0 || h->device >= kMaxDevices) { h->err = "mpcb_step_static: device index out of range"; return MPCB_EINVAL; }
    if (!g_step_scratch_ptr[h->device]) {
        int cur = -1;
        CUDA_TRY(h, cudaGetDevice(&cur));
        if (cur != h->device) CUDA_TRY(h, cudaSetDevice(h->device));
        int* ptr = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_step_scratch);
        if (cur != h->device) cudaSetDevice(cur);
        if (e != cudaSuccess || !ptr) {
            h->err = std::string("mpcb_step_static: cudaGetSymbolAddress(g_step_scratch): ") + cudaGetErrorString(e);
            return MPCB_ECUDA;
        }
        g_step_scratch_ptr[h->device] = ptr;
    }
    *out = g_step_scratch_ptr[h->device];
    return MPCB_OK;
}
}  // namespace

extern "C" int mpcb_step_static(mpcb_handle* h, int B, const double* state_dev, const int* route_dev, const int* ncourse_dev,
                                const double* L_dev, int* target_ind_dev, int* warm_dev, double* oa_dev, double* od_dev,
                                double* traj_dev, double* xref_dev, double* di_ai_dev, double* xdev_dev, int* status_dev,
                                int* iters_dev, void* stream) {
    if (!h) return MPCB_EINVAL;
    if (B <= 0 || !state_dev || !route_dev || !ncourse_dev || !target_ind_dev || !warm_dev ||
        !oa_dev || !od_dev || !di_ai_dev || !status_dev) { h->err = "mpcb_step_static: bad arguments"; return MPCB_EINVAL; }
    if (B > MPCB_STATIC_MAX_BATCH || B > h->max_batch) {
        h->err = "mpcb_step_static: B=" + std::to_string(B) + " exceeds the static scratch (MPCB_STATIC_MAX_BATCH=" +
                 std::to_string(MPCB_STATIC_MAX_BATCH) + ") or the handle's max_batch; use mpcb_step";
        return MPCB_EINVAL;
    }
    if (h->cfg.T > 31) { h->err = "mpcb_step_static: the long-horizon kernel (T > 31) uses no scratch; use mpcb_step"; return MPCB_EINVAL; }
    if (!h->courses_dev) { h->err = "mpcb_step_static: no courses set"; return MPCB_EINVAL; }
    int* scr = nullptr;
    int rc = static_scratch_ptr(h, &scr);
    if (rc != MPCB_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    static_scratch_prepare<<<1, 1024, 0, st>>>(scr, B);
    CUDA_TRY(h, cudaGetLastError());
    h->launches++;
    StepParams p = make_params(h);
    p.B = B; p.state = state_dev; p.route = route_dev; p.ncourse = ncourse_dev; p.L = L_dev;
    p.target_ind = target_ind_dev; p.warm = warm_dev; p.oa = oa_dev; p.od = od_dev; p.traj = traj_dev;
    p.xref = xref_dev; p.di_ai = di_ai_dev; p.xdev = xdev_dev; p.status = status_dev; p.iters = iters_dev;
    p.lpt_hist = scr; p.lpt_order = scr + kStaticScratchHead; p.lpt_rank = scr + kStaticScratchHead + MPCB_STATIC_MAX_BATCH;
    return dispatch_step(h, p, st);
}

#ifdef MPCB_TRACE
extern "C" void mpcb_debug_set_trace(double* trace_dev) { g_trace_dev = trace_dev; }
#endif
