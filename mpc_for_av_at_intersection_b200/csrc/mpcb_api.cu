// mpcb_api.cu -- C ABI of libmpcb.so (see include/mpcb.h) and the small A8 kernels.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/mpcb.h"
#include "mpcb_step.cuh"

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

constexpr int kHorizons[] = {10, 13, 20};

template <int T, int WARPS>
int launch_step(mpcb_handle* h, const StepParams& p, cudaStream_t st) {
    const size_t smem = sizeof(mpcb::WarpSmem<T>) * WARPS;
    auto kern = mpcb::mpc_step_kernel<T, WARPS>;
    static thread_local int configured_dev = -1;
    if (configured_dev != h->device) {
        CUDA_TRY(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured_dev = h->device;
    }
    const int grid = (p.B + WARPS - 1) / WARPS;
    kern<<<grid, WARPS * 32, smem, st>>>(p);
    CUDA_TRY(h, cudaGetLastError());
    h->launches++;
    return MPCB_OK;
}

StepParams make_params(const mpcb_handle* h) {
    const mpcb_config& c = h->cfg;
    StepParams p;
    memset(&p, 0, sizeof(p));
    p.R = h->R; p.ncmax = h->ncmax; p.courses = h->courses_dev; p.dl = h->dl_dev;
    p.w_perp = c.w_perp; p.w_para = c.w_para; p.R0 = c.R[0]; p.R1 = c.R[1]; p.Rd0 = c.Rd[0]; p.Rd1 = c.Rd[1];
    p.Qv = c.Q_v_yaw[0]; p.Qyaw = c.Q_v_yaw[1];
    p.Qf0 = c.Qf[0] * c.T; p.Qf1 = c.Qf[1] * c.T; p.Qf2 = c.Qf[2] * c.T; p.Qf3 = c.Qf[3] * c.T;   // mpc.py:25
    p.dsteer_dt = (c.max_dsteer_deg * (3.141592653589793 / 180.0)) * c.dt;                        // mpc.py:34,184
    p.max_accel = c.max_accel; p.max_decel = c.max_decel; p.max_steer = c.max_steer;
    p.max_speed = c.max_speed; p.min_speed = c.min_speed; p.r_end = c.r_end; p.min_ref_speed = c.min_ref_speed;
    p.dt = c.dt; p.wheelbase = c.wheelbase; p.max_iter = c.max_iter; p.ipm_max_iter = c.ipm_max_iter;
    p.tol_mu = c.ipm_tol_mu; p.tol_p = c.ipm_tol_p; p.tol_d = c.ipm_tol_d;
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
    c->ipm_max_iter = 40; c->ipm_tol_mu = 1e-12; c->ipm_tol_p = 1e-10; c->ipm_tol_d = 1e-9;
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
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0 || device < 0 || device >= n) {
        g_create_error = "mpcb_create: no usable CUDA device (this library has no CPU fallback)";
        return MPCB_ENODEV;
    }
    mpcb_handle* h = new mpcb_handle();
    h->cfg = *cfg; h->device = device; h->max_batch = max_batch;
    if (cudaSetDevice(device) != cudaSuccess) { g_create_error = "cudaSetDevice failed"; delete h; return MPCB_ECUDA; }
    cudaDeviceGetAttribute(&h->smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device);
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
    if (!h || !courses_host || !dl_host || R <= 0 || ncmax <= 0) { if (h) h->err = "mpcb_set_courses_host: bad arguments"; return MPCB_EINVAL; }
    (void)len_host;
    CUDA_TRY(h, cudaSetDevice(h->device));
    if (h->courses_dev) { cudaFree(h->courses_dev); h->courses_dev = nullptr; }
    if (h->dl_dev) { cudaFree(h->dl_dev); h->dl_dev = nullptr; }
    const size_t bytes = (size_t)R * 3 * ncmax * sizeof(double);
    CUDA_TRY(h, cudaMalloc(&h->courses_dev, bytes));
    CUDA_TRY(h, cudaMalloc(&h->dl_dev, R * sizeof(double)));
    CUDA_TRY(h, cudaMemcpy(h->courses_dev, courses_host, bytes, cudaMemcpyHostToDevice));
    CUDA_TRY(h, cudaMemcpy(h->dl_dev, dl_host, R * sizeof(double), cudaMemcpyHostToDevice));
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
    switch (h->cfg.T) {
        case 10: return launch_step<10, 4>(h, p, st);
        case 13: return launch_step<13, 4>(h, p, st);
        case 20: return launch_step<20, 4>(h, p, st);
        default: h->err = "mpcb_step: unsupported horizon"; return MPCB_EINVAL;
    }
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
