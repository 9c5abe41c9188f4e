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
#define MPCB_ST_INACCURATE 1  /* THIS solver's reduced residual class (dual 1e-6 x gradient scale, primal 1e-7, gap 1e-7 x
                                 max(1, 1e-4 gradient scale) -- ECOS too judges the gap relative to the objective -- on
                                 a verified iterate), accepted like cvxpy.OPTIMAL_INACCURATE is (ibid.).  It is an analogue, not
                                 ECOS's own class: cvxpy -> ECOS is absent from the reference tree and unpinned, so which
                                 problems ECOS would call inaccurate cannot be reproduced.  OPTIMAL (0) means the configured
                                 tolerances or at least the 1e-8 class on a verified iterate.                          */
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
    int ipm_max_iter;      /* default 100                                                    */
    double ipm_tol_mu;     /* complementarity gap target, default 1e-12                      */
    double ipm_tol_p;      /* primal residual,             default 1e-10                     */
    double ipm_tol_d;      /* dual residual (absolute part), default 1e-9                    */
    /* inter-vehicle coupling (row A9): CarDimensions.bounding_box_size (main/lib/car_dimensions.py:82-107) and the
     * literals of main/scenarios/mpc_intersection.py:81-84 */
    double car_width;      /* 2.0 (bicycle model) / 2.04 s (Prius): circle radius = width / sqrt(2)  */
    double car_length;     /* wheelbase + 0.64 = 3.5 / 4.84 s                                        */
    double coupling_horizon; /* TIME_HORIZON = 7.0 s -> 35 predicted obstacle frames at dt = 0.2       */
    int frame_window;      /* FRAME_WINDOW = 20: obstacle trajectories are time-shifted by -20..+20 frames */
    /* controller variant main/lib/mpc_with_speed.py: speed reference xref[2,t] = cv[idx_t] (:104) with cv = ref_speed
     * (its MAX_SPEED = 25/3.6, :36,280) and 0 from the per-problem cutoff index on (:281-282).  0 = main/lib/mpc.py */
    double ref_speed;
    /* controller variant main/lib/mpc_jerk.py: NX = 5, the fifth state integrates the acceleration input from a FREE initial
     * value and feeds the speed (`A[2,4] = dt`, `A[4,4] = 1`, `B[4,0] = dt`, :61-86), plus a penalty on its increments
     * (jerk_penalty_weight, :30,190).  jerk_model = 1 selects the jerk kernel (T = 10, 13, 20); 0 = main/lib/mpc.py */
    int jerk_model;
    double jerk_weight;
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
/* horizons this build has kernels for (10, 13, 20: one warp per problem; 40, 80: one CTA per problem, the sizes of
 * BASELINE.json configs[4]); returns count, fills up to cap entries */
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

/* ---- allocation-free step ---------------------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arguments, results and
 * asynchronous behaviour as mpcb_step, plus the longest-first launch order that otherwise needs a session: the order's
 * scratch (iteration histogram, per-problem ranks, slot -> problem table) lives in a file-scope static __device__ array of
 * 129 KB that every GPU loading the library holds until exit.  The call allocates nothing (no cudaMalloc / cudaFree, no
 * allocating library routine); the first call on a device looks the array's address up (cudaGetSymbolAddress), so make
 * that first call outside a capture.
 *   - B > MPCB_STATIC_MAX_BATCH (or > the handle's max_batch) and horizons T > 31 (their kernel has no scratch) are
 *     rejected with MPCB_EINVAL and a message; use mpcb_step there.
 *   - All calls on a device share the array, whatever the handle.  It keeps the last call's slot -> problem table (reset
 *     when B changes; the histogram is cleared by every call); the table never changes a result, only the launch order.
 *   - Calls that overlap on the GPU (two streams, two handles, two threads) race on the array: serialise them. */
#define MPCB_STATIC_MAX_BATCH 16384
int mpcb_step_static(mpcb_handle* h, int B, const double* state_dev, const int* route_dev, const int* ncourse_dev,
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

/* ---- resident session -------------------------------------------------------------------------
 * The reference keeps one `MPC` object per vehicle alive across the closed loop
 * (main/scenarios/mpc_intersection.py:71,95-159) and mutates it in MPC.step (main/lib/mpc.py:289-292).
 * A session is the [B,...] device-resident equivalent owned by the handle: per control step only the
 * vehicle states go in and the applied controls come out; everything else (target_ind, warm start,
 * ox/oy/ov/oyaw, xref) stays in HBM and is fetched on demand. */

/* fields of the session (mpcb_session_fetch_host / _store_host / _device_ptr) */
#define MPCB_F_STATE 0    /* double [B,4]      x, y, yaw, v                                  */
#define MPCB_F_OA 1       /* double [B,T]      mpc.oa                                        */
#define MPCB_F_OD 2       /* double [B,T]      mpc.odelta                                    */
#define MPCB_F_TRAJ 3     /* double [B,4,T+1]  mpc.ox, oy, ov, oyaw                          */
#define MPCB_F_XREF 4     /* double [B,4,T+1]  mpc.xref                                      */
#define MPCB_F_DI_AI 5    /* double [B,2]      mpc.di, mpc.ai                                */
#define MPCB_F_XDEV 6     /* double [B]        get_current_xref_deviation()                  */
#define MPCB_F_TARGET 7   /* int    [B]        mpc.target_ind                                */
#define MPCB_F_WARM 8     /* int    [B]        mpc.oa is not None                            */
#define MPCB_F_STATUS 9   /* int    [B]        MPCB_ST_*                                     */
#define MPCB_F_ITERS 10   /* int    [B]        interior-point iterations                     */
#define MPCB_F_DONE 11    /* int    [B]        closed loop: 0 running, 1 is_goal, 2 dead (MPCB_ST_NEAREST) */
#define MPCB_F_NCOURSE 12 /* int    [B]        len(mpc.cx)                                   */
#define MPCB_F_ROUTE 13   /* int    [B]                                                      */
#define MPCB_F_GOAL 14    /* double [B,2]      mpc.goal                                      */

/* replaces: constructing B `MPC` objects (main/lib/mpc.py:243-271): target_ind = 0 (or
 * target_ind_host), oa = odelta = None, di = ai = 0, goal = end of the full course.
 * ncourse_host NULL = full course lengths; L_host NULL = cfg.wheelbase.  Synchronous. */
int mpcb_session_begin(mpcb_handle* h, int B, const int* route_host, const int* ncourse_host,
                       const double* L_host, const int* target_ind_host);
/* replaces: MPC.set_trajectory_fromarray(full[:cutoff]) (main/lib/mpc.py:275-278) for the batch */
int mpcb_session_set_ncourse_host(mpcb_handle* h, const int* ncourse_host);
/* replaces: main/lib/mpc_sensitivity.py:150-158, which re-reads w_perp, w_para, R, Rd, Q_v_yaw, Qf from its JSON on
 * every solve so that main/scenarios/mpc_sensitivity_analysis_comulative.py:152-157 can sweep them: here every problem
 * of the batch carries its own row, so a whole sweep (5 values x 6 parameters x routes) is ONE batch.
 * weights_host [B,12]: w_perp, w_para, R[0], R[1], Rd[0], Rd[1], Q_v_yaw[0], Q_v_yaw[1], Qf[0..3] (as in the file:
 * the library scales Qf by T).  NULL = back to the handle's config.  Synchronous. */
int mpcb_session_set_weights_host(mpcb_handle* h, const double* weights_host);
/* replaces: the cutoff_idx argument of mpc_with_speed.MPC.set_trajectory_fromarray (main/lib/mpc_with_speed.py:276-282):
 * the speed reference cv is cfg.ref_speed up to cutoff_idx_host[b] and 0 from there on (999 or any index past the
 * course = no cut).  NULL = no cut anywhere.  Synchronous. */
int mpcb_session_set_speed_cutoff_host(mpcb_handle* h, const int* cutoff_idx_host);
/* replaces: `di, ai = mpc.step(state)` for the batch with HOST buffers: copies state[B,4] in, runs the
 * step kernel on the resident session, copies di_ai[B,2] (steer, accel) and status[B] (may be NULL)
 * out.  Synchronous. */
int mpcb_session_step_host(mpcb_handle* h, const double* state_host, double* di_ai_host, int* status_host);
/* replaces: `di, ai = mpc.step(state)` TOGETHER with the reads a scenario script does afterwards (mpc.ox, mpc.oy, mpc.xref,
 * mpc.target_ind, mpc.oa ..., main/lib/mpc.py:289-292, main/scenarios/mpc_intersection.py:297-302) in one synchronisation:
 * optional target_ind / warm overrides in (NULL = keep the session's), state in, the step, every non-NULL output out.
 * This is what the B = 1 facade compat.mpc.MPC.step calls.  Synchronous. */
int mpcb_session_step_all_host(mpcb_handle* h, const double* state_host, const int* target_ind_host, const int* warm_host,
                               double* di_ai_host, int* status_host, int* target_ind_out, int* iters_out, double* oa_out,
                               double* od_out, double* traj_out, double* xref_out);
/* same step on the state already resident in MPCB_F_STATE; asynchronous on `stream`, capturable */
int mpcb_session_step_dev(mpcb_handle* h, void* stream);
/* replaces: the loop body of main/scenarios/mpc_intersection.py:95-159 without obstacles, nsteps
 * times: is_goal -> break (done=1) | mpc.step | simulation.step, one kernel launch per control step.
 * history_dev NULL or double [nsteps,B,8] (x, y, yaw, v after the plant step, delta, a,
 * xref_deviation, status) -- the columns of lib/simulation.py History; rows of finished problems are
 * left untouched.  flags: MPCB_CL_NO_GOAL_TEST keeps every problem solving on every step (throughput
 * measurement: no work is skipped).  Asynchronous on `stream`, capturable. */
#define MPCB_CL_NO_GOAL_TEST 1
int mpcb_session_closed_loop_dev(mpcb_handle* h, int nsteps, double* history_dev, int flags, void* stream);
void* mpcb_session_device_ptr(mpcb_handle* h, int what);
long long mpcb_session_field_bytes(const mpcb_handle* h, int what);
int mpcb_session_fetch_host(mpcb_handle* h, int what, void* dst, long long bytes);
int mpcb_session_store_host(mpcb_handle* h, int what, const void* src, long long bytes);

/* ---- inter-vehicle coupling (row A9) ---------------------------------------------------------------
 * replaces, per ego vehicle and control step, main/scenarios/mpc_intersection.py:100-139:
 * calc_nearest_index_in_direction on the full path -> resample_curve ego prediction
 * (main/lib/trajectories.py:58-86) -> MovingObstaclesPrediction.state_prediction
 * (main/lib/moving_obstacles_prediction.py:21-47) -> check_collision_moving_cars
 * (main/lib/collision_avoidance.py:66-104) -> get_cutoff_curve_by_position_idx (:107-119) -> cutoff_idx.
 *   state_dev     [B,4]  x, y, yaw, v of every ego
 *   route_dev     [B]
 *   obs_dev       [B,MPCB_MAX_OBSTACLES,6] obstacle states in MovingObstacle*.get() order (x, y, v, yaw, a, steer);
 *                 ignored (may be NULL) when vehicles_per_scenario > 0
 *   n_obs_dev     [B]    obstacles per ego (<= MPCB_MAX_OBSTACLES), NULL = 0
 *   done_dev      [B] or NULL: finished egos are skipped (their cut is kept); in/out: an ego whose nearest-index lookup
 *                 hits the reference's Exception("something wrong") (main/lib/trajectories.py:120) is marked 2 (dead)
 *   agent_idx_dev [B]    in/out  traj_agent_idx
 *   ncourse_dev   [B]    in/out  in: len(tmp_trajectory) of the previous step (0 = None), out: new cutoff_idx
 *                        (= the full course length when nothing collides) -- feed it to mpcb_step as ncourse
 *   collided_dev  [B] out (0/1), coll_idx_dev [B] out (index into full[agent_idx:], -1), coll_xy_dev [B,2] out,
 *   n_res_dev     [B] out len(trajectory_res), flags_dev [B] out (bit0 nearest-index exception, bit1 overflow);
 *                 any of the five may be NULL
 *   vehicles_per_scenario  > 0: problems are grouped in scenarios of that many consecutive egos and every ego takes
 *                 the OTHER egos of its scenario as obstacles (constant speed, zero steer: the intent stated at
 *                 main/scenarios/interactive_mpc_temp.py:168); all egos read start-of-step states (Jacobi order,
 *                 ibid.:124-210).  0: obstacles come from obs_dev. */
#define MPCB_MAX_OBSTACLES 8
int mpcb_coupling(mpcb_handle* h, int B, const double* state_dev, const int* route_dev, const double* obs_dev,
                  const int* n_obs_dev, int* done_dev, int* agent_idx_dev, int* ncourse_dev, int* collided_dev,
                  int* coll_idx_dev, double* coll_xy_dev, int* n_res_dev, int* flags_dev, int vehicles_per_scenario,
                  void* stream);
/* same with HOST buffers (synchronous); in/out arrays are updated in place */
int mpcb_coupling_host(mpcb_handle* h, int B, const double* state, const int* route, const double* obs,
                       const int* n_obs, int* agent_idx, int* ncourse, int* collided, int* coll_idx, double* coll_xy,
                       int* n_res, int* flags, int vehicles_per_scenario);
/* replaces: the multi-vehicle closed loop (SURVEY.md section 8a, A9 spec; main/scenarios/interactive_mpc_temp.py:124-210
 * as intended): nsteps x (is_goal | coupling against the other egos of the scenario | mpc.step | simulation.step),
 * two kernel launches per control step on the resident session.  history_dev as mpcb_session_closed_loop_dev. */
int mpcb_session_multi_agent_dev(mpcb_handle* h, int nsteps, int vehicles_per_scenario, double* history_dev, int flags,
                                 void* stream);
#define MPCB_F_AGENT_IDX 15 /* int [B] traj_agent_idx */
#define MPCB_F_COLLIDED 16  /* int [B] */
#define MPCB_F_COLL_IDX 17  /* int [B] */

/* ---- scripted cross traffic on the device ------------------------------------------
 * replaces: the `MovingObstacleTIntersection` objects of the reference scenario (main/lib/moving_obstacles.py:165-231,
 * constructed at main/scenarios/mpc_intersection.py:41-44) and their part of the loop body -- `o.get()` feeding the
 * predictions (:118-121) and `o.step()` after the solve (:154-156).
 *   script_host [B, MPCB_MAX_OBSTACLES, 4]  constructor arguments per obstacle: direction (>= 0: from the left at
 *               (-30, -3) heading 0; < 0: from the right at (30, 3) heading pi), turning (0/1), speed, offset
 *               (seconds before it starts moving; <= 0 = None).  dt and the wheelbase are the handle's.
 *   n_obs_host  [B] obstacles of each ego's scenario (0..MPCB_MAX_OBSTACLES)
 * Resets every script to its constructor state (counter = 0).  Synchronous. */
int mpcb_session_set_scripted_host(mpcb_handle* h, const double* script_host, const int* n_obs_host);
/* replaces: the whole loop body of main/scenarios/mpc_intersection.py:95-159 WITH its scripted obstacles, nsteps times,
 * with no host round trip: per control step one coupling launch (is_goal | obstacle get() + step() | nearest index |
 * predictions | first collision | path cut) and one step launch (mpc.step | simulation.step | history row).
 * history_dev as mpcb_session_closed_loop_dev; obs_history_dev NULL or double [nsteps,B,MPCB_MAX_OBSTACLES,6]: the
 * rows `o.get()` returned at every step (the reference's `obstacles_positions`, :92,155).  Asynchronous, capturable. */
int mpcb_session_scripted_loop_dev(mpcb_handle* h, int nsteps, double* history_dev, double* obs_history_dev, int flags,
                                   void* stream);
#define MPCB_F_SCRIPT 18    /* double [B,MPCB_MAX_OBSTACLES,8] xc, yc, theta, counter, direction, turning, speed, offset */
#define MPCB_F_OBS 19       /* double [B,MPCB_MAX_OBSTACLES,6] last get() rows (when no obs_history_dev was given) */
#define MPCB_F_NOBS 20      /* int    [B] scripted obstacles per ego */
#define MPCB_F_LAUNCH_ORDER 21 /* int [B] launch slot -> problem of the next step: longest-first by the interior-point
                                 iterations of the last step (a permutation; identity until the first step and whenever the
                                 batch is under one or over a dozen waves of resident warps).  No effect on any result. */

/* ---- motion-primitive A* planner ------------------------------------------------------
 * replaces: `MotionPrimitiveSearch(scenario, car_dimensions, mps, margin).run()` (main/lib/motion_primitive_search_modified.py:
 * 18-58) = the A* loop of main/lib/a_star.py:31-78 over the neighbour function :101-149 (collision check of every
 * primitive's circle-centre points against the obstacles' half-planes, main/lib/obstacles.py:157-176), the goal test
 * and heuristic :64-86 and the trajectory assembly :151-165 -- for B searches at once, one warp per search.
 * All pointers are HOST memory; synchronous; no handle (the planner runs once per vehicle, before the MPC loop).
 *   start, goal      [B,3]  (x, y, theta): Scenario.start, Scenario.goal_point (main/lib/scenario.py:7-13)
 *   goal_box         [B,4]  x1, y1, x2, y2 of Scenario.goal_area (a BoxObstacle, obstacles.py:46-58)
 *   allowed          [B]    Scenario.allowed_goal_theta_difference
 *   obs_set          [B]    which obstacle set each search uses (0..S-1)
 *   halfplanes       [S,32,8,3], n_halfplanes [S,32], n_obstacles [S]: Obstacle.to_convex(margin) rows (a, b, c) --
 *                    a point is inside an obstacle where a x + b y + c <= 0 for every row (obstacles.py:96-105,135-151)
 *   mp_points        [P,NP,3] primitive points (x, y, theta) in the primitive's start frame, mp_n_points [P] valid points,
 *                    mp_length [P] MotionPrimitive.total_length; P <= 16, in the order the reference iterates `mps`
 *   mp_cc_points     [P,NCC,2] circle-centre collision points of every primitive (:35-52), NCC <= 16
 *   max_open, max_closed, max_path: capacities of the open list, the closed set and the path (nodes) per search
 * Outputs: status [B] MPCB_PLAN_*, cost [B] (NaN unless OK), n_path [B] nodes on the path, mp_seq [B,max_path]
 * primitive index per edge (-1 padded), nodes [B,max_path,3], n_expanded [B] (nodes taken off the queue and kept,
 * = len(debug_data)), n_pushed [B]; traj [B,(max_path-1)*(NP-1),3] + n_traj [B] or NULL. */
#define MPCB_PLAN_OK 0
#define MPCB_PLAN_NO_SOLUTION 1   /* Exception("No solution found.") (main/lib/a_star.py:78) */
#define MPCB_PLAN_OVERFLOW 2      /* a capacity was too small: enlarge max_open / max_closed / max_path */
int mpcb_plan_routes_host(int device, int B, const double* start, const double* goal, const double* goal_box,
                          const double* allowed, const int* obs_set, int S, const double* halfplanes, const int* n_halfplanes,
                          const int* n_obstacles, int P, int NP, const double* mp_points, const int* mp_n_points,
                          const double* mp_length, int NCC, const double* mp_cc_points, int max_open, int max_closed,
                          int max_path, int* status, double* cost, int* n_path, int* mp_seq, double* nodes, int* n_expanded,
                          int* n_pushed, double* traj, int* n_traj);
const char* mpcb_plan_last_error(void);

/* kernels launched by this handle so far (bench.py's gpu_launches) */
long long mpcb_launch_count(const mpcb_handle* h);

/* FP64 FMA peak of `device` measured by a register-resident FMA chain kernel (TFLOP/s); the
 * roofline denominator for the solve kernel (MEASURED_PEAKS.json has no FP64 figure). */
int mpcb_measure_fp64_peak(int device, double* tflops_out);

#ifdef __cplusplus
}
#endif
#endif /* MPCB_H */
