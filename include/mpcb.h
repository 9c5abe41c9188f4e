/*
 * mpcb.h -- C ABI of libmpcb.so: batched kinematic-bicycle MPC for B200 (sm_100a).
 *
 * The reference (SaeedRahmani/MPC_for_AV_at_Intersection) has no FFI layer: its
 * boundary for this path is the Python class `MPC` in main/lib/mpc.py:242-326 and
 * the JSON file main/config/mpc_config.json.  Each entry point below names the
 * reference interface it replaces.  All signatures are plain C (pointers and
 * sizes); nothing from torch or C++ crosses this boundary.
 *
 * Conventions
 *   - `*_dev` pointers are caller-owned DEVICE memory (e.g. torch CUDA tensors),
 *     float64 / int32, contiguous, layouts as commented.
 *   - `*_host` entry points take HOST pointers and do the host<->device copies
 *     themselves through the handle's pinned staging buffers.
 *   - Device entry points are asynchronous on `stream` (a cudaStream_t passed as
 *     void*, NULL = legacy default stream), never synchronise, and are CUDA-graph
 *     capturable.  A handle is not thread-safe; distinct handles are independent.
 *   - Return value: 0 on success, negative MPCB_E* on API errors (bad sizes, CUDA
 *     failure); mpcb_last_error() gives the text.  Per-problem numerical outcomes
 *     are reported in status[B], never as an error return.
 *   - There is no CPU fallback: without a CUDA device mpcb_create fails.
 */
#ifndef MPCB_H
#define MPCB_H

#ifdef __cplusplus
extern "C" {
#endif

#define MPCB_VERSION 100

/* error codes */
#define MPCB_OK 0
#define MPCB_EINVAL (-1)      /* bad argument / unsupported horizon            */
#define MPCB_ECUDA (-2)       /* CUDA runtime failure                           */
#define MPCB_ENODEV (-3)      /* no usable CUDA device                          */
#define MPCB_ENOMEM (-4)

/* per-problem status written to status[B] */
#define MPCB_ST_OPTIMAL 0     /* cvxpy.OPTIMAL            (main/lib/mpc.py:196)  */
#define MPCB_ST_INACCURATE 1  /* cvxpy.OPTIMAL_INACCURATE (accepted, ibid.)      */
#define MPCB_ST_FAILED 2      /* solve failed: ai = MAX_DECEL, di kept, warm start dropped (mpc.py:204-208,294-297) */
#define MPCB_ST_NEAREST 3     /* Exception("something wrong") in calc_nearest_index_in_direction
                                 (main/lib/trajectories.py:120): the problem's state is left untouched */

/* POD mirror of main/config/mpc_config.json plus the constants main/lib/mpc.py:17-36 and
 * main/lib/simulation.py:23-25 derive or hard-code.  Fill with mpcb_default_config() first. */
typedef struct mpcb_config {
    int T;                 /* horizon "T" (default 13; BASELINE configs use 20)              */
    int max_iter;          /* "MAX_ITER": linearise+solve passes per step (default 1)        */
    double w_perp, w_para; /* "w_perp", "w_para"                                             */
    double R[2];           /* "R"  input cost (accel, steer)                                 */
    double Rd[2];          /* "Rd" input-difference cost                                     */
    double Q_v_yaw[2];     /* "Q_v_yaw"                                                      */
    double Qf[4];          /* "Qf" as in the file (UNscaled; the library multiplies by T as mpc.py:25 does) */
    double goal_dis;       /* "GOAL_DIS"                                                     */
    double stop_speed;     /* "STOP_SPEED"                                                   */
    double max_dsteer_deg; /* "MAX_DSTEER" in deg/s as in the file (mpc.py:34 converts)      */
    double max_accel;      /* "MAX_ACCEL"                                                    */
    double max_decel;      /* "MAX_DECEL"                                                    */
    double max_steer;      /* Simulation.MAX_STEER  = deg2rad(45)   (simulation.py:23)       */
    double max_speed;      /* Simulation.MAX_SPEED  = 30/3.6        (simulation.py:24)       */
    double min_speed;      /* Simulation.MIN_SPEED  = -5            (simulation.py:25)       */
    double r_end;          /* input cost once the reference reached the path end: 10 (mpc.py:178) */
    double min_ref_speed;  /* floor of the reference speed: 10/3.6 (mpc.py:96)               */
    double dt;             /* sample time (MPC.__init__ dt, default 0.2)                     */
    double wheelbase;      /* CarDimensions.distance_back_to_front_wheel used when L_dev==NULL (2.86) */
    /* interior-point controls (no reference counterpart: cvxpy->ECOS defaults are abstol=reltol=feastol=1e-8) */
    int ipm_max_iter;      /* default 40                                                     */
    double ipm_tol_mu;     /* complementarity gap target, default 1e-12                      */
    double ipm_tol_p;      /* primal residual,             default 1e-10                     */
    double ipm_tol_d;      /* dual residual (absolute part), default 1e-9                    */
} mpcb_config;

typedef struct mpcb_handle mpcb_handle;

int mpcb_version(void);
void mpcb_default_config(mpcb_config* cfg);
/* number of CUDA devices visible, or a negative error */
int mpcb_device_count(void);

/* replaces: `MPC(...)` construction + module-level config load (main/lib/mpc.py:13-36,243-271).
 * Allocates the handle's workspace on `device` for at most `max_batch` problems. */
int mpcb_create(const mpcb_config* cfg, int max_batch, int device, mpcb_handle** out);
void mpcb_destroy(mpcb_handle* h);
const char* mpcb_last_error(const mpcb_handle* h);   /* h may be NULL: last create error */
/* horizons this build has kernels for; returns count, fills up to cap entries */
int mpcb_supported_horizons(int* out, int cap);

/* replaces: smooth_yaw (main/lib/mpc.py:43-55).  Host, in place, returns yaw for chaining parity. */
void mpcb_smooth_yaw(double* yaw, int n);

/* replaces: the cx, cy, cyaw, dl arguments of MPC.__init__ / set_trajectory_fromarray
 * (main/lib/mpc.py:243-278).  courses_host is [R][3][ncmax] (x row, y row, yaw row per route, yaw
 * ALREADY smoothed), len_host[R] the full lengths, dl_host[R] the course tick.  Copies to the device. */
int mpcb_set_courses_host(mpcb_handle* h, const double* courses_host, const int* len_host,
                          const double* dl_host, int R, int ncmax);

/* replaces: MPC.step (main/lib/mpc.py:280-299) for B independent problems.
 *   state_dev      [B,4]  x, y, yaw, v            (State field order, simulation.py:12-19)
 *   route_dev      [B]    course id of each problem
 *   ncourse_dev    [B]    current (possibly cut) course length = len(mpc.cx)
 *   L_dev          [B] or NULL (then cfg.wheelbase)
 *   target_ind_dev [B]    in/out  mpc.target_ind
 *   warm_dev       [B]    in/out  1 if oa/od hold a previous solution (0 <=> `oa is None`)
 *   oa_dev, od_dev [B,T]  in/out  mpc.oa, mpc.odelta (UNshifted warm start in, solution out)
 *   traj_dev       [B,4,T+1] out  ox, oy, ov, oyaw
 *   xref_dev       [B,4,T+1] out  mpc.xref
 *   di_ai_dev      [B,2]  in/out  mpc.di, mpc.ai  (the applied controls: steer, accel)
 *   xdev_dev       [B]    out     get_current_xref_deviation() (mpc.py:301-308); NaN if undefined
 *   status_dev     [B]    out     MPCB_ST_*
 *   iters_dev      [B]    out     interior-point iterations of the last pass
 */
int mpcb_step(mpcb_handle* h, int B, const double* state_dev, const int* route_dev, const int* ncourse_dev,
              const double* L_dev, int* target_ind_dev, int* warm_dev, double* oa_dev, double* od_dev,
              double* traj_dev, double* xref_dev, double* di_ai_dev, double* xdev_dev, int* status_dev,
              int* iters_dev, void* stream);

/* Same call with HOST buffers (any pointer except state/route/ncourse may be NULL to skip that
 * output; in/out arrays are read and written).  Synchronous. */
int mpcb_step_host(mpcb_handle* h, int B, const double* state, const int* route, const int* ncourse,
                   const double* L, int* target_ind, int* warm, double* oa, double* od, double* traj,
                   double* xref, double* di_ai, double* xdev, int* status, int* iters);

/* replaces: Simulation.step / Bicycle.step (main/lib/simulation.py:35-47, main/bicycle/main.py:28-41).
 * state_dev [B,4] in/out; di_ai_dev [B,2] (steer, accel); active_dev [B] or NULL (0 = frozen). */
int mpcb_plant_step(mpcb_handle* h, int B, double* state_dev, const double* di_ai_dev, const double* L_dev,
                    const int* active_dev, void* stream);

/* replaces: MPC.is_goal (main/lib/mpc.py:310-326).  goal_dev [B,2]; out_dev [B] 0/1. */
int mpcb_is_goal(mpcb_handle* h, int B, const double* state_dev, const double* goal_dev,
                 const int* target_ind_dev, const int* ncourse_dev, int* out_dev, void* stream);

/* kernels launched by this handle so far (bench.py's gpu_launches) */
long long mpcb_launch_count(const mpcb_handle* h);

/* FP64 FMA peak of `device` measured by a register-resident FMA chain kernel (TFLOP/s); the
 * roofline denominator for the solve kernel (MEASURED_PEAKS.json has no FP64 figure). */
int mpcb_measure_fp64_peak(int device, double* tflops_out);

#ifdef __cplusplus
}
#endif
#endif /* MPCB_H */
