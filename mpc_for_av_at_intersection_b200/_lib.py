"""ctypes binding of libmpcb.so (include/mpcb.h).  There is no fallback: if the shared library is
missing or no CUDA device is usable, the product raises."""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MPCB_LIB_PATH") or os.path.join(_HERE, "libmpcb.so")     # override: A/B runs of kernel variants
CSRC = os.path.join(_HERE, "csrc")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-shared", "-Xcompiler", "-fPIC",
              "-cudart", "shared"]      # libcudart.so.12 of the image / of torch: no statically linked runtime inside the artefact


class MpcbError(RuntimeError):
    pass


class mpcb_config(C.Structure):
    _fields_ = [
        ("T", C.c_int), ("max_iter", C.c_int),
        ("w_perp", C.c_double), ("w_para", C.c_double),
        ("R", C.c_double * 2), ("Rd", C.c_double * 2), ("Q_v_yaw", C.c_double * 2), ("Qf", C.c_double * 4),
        ("goal_dis", C.c_double), ("stop_speed", C.c_double), ("max_dsteer_deg", C.c_double),
        ("max_accel", C.c_double), ("max_decel", C.c_double), ("max_steer", C.c_double),
        ("max_speed", C.c_double), ("min_speed", C.c_double), ("r_end", C.c_double),
        ("min_ref_speed", C.c_double), ("dt", C.c_double), ("wheelbase", C.c_double),
        ("ipm_max_iter", C.c_int), ("ipm_tol_mu", C.c_double), ("ipm_tol_p", C.c_double),
        ("ipm_tol_d", C.c_double),
        ("car_width", C.c_double), ("car_length", C.c_double), ("coupling_horizon", C.c_double),
        ("frame_window", C.c_int),
        ("ref_speed", C.c_double),
        ("jerk_model", C.c_int), ("jerk_weight", C.c_double),
    ]


EXPORTS = [
    "mpcb_version", "mpcb_default_config", "mpcb_device_count", "mpcb_create", "mpcb_destroy",
    "mpcb_last_error", "mpcb_supported_horizons", "mpcb_smooth_yaw", "mpcb_set_courses_host", "mpcb_step",
    "mpcb_step_host", "mpcb_plant_step", "mpcb_is_goal", "mpcb_launch_count", "mpcb_measure_fp64_peak",
    "mpcb_session_begin", "mpcb_session_set_ncourse_host", "mpcb_session_step_host", "mpcb_session_step_dev",
    "mpcb_session_closed_loop_dev", "mpcb_session_device_ptr", "mpcb_session_field_bytes",
    "mpcb_session_fetch_host", "mpcb_session_store_host",
    "mpcb_coupling", "mpcb_coupling_host", "mpcb_session_multi_agent_dev",
    "mpcb_session_set_scripted_host", "mpcb_session_scripted_loop_dev",
    "mpcb_session_set_weights_host", "mpcb_session_set_speed_cutoff_host",
    "mpcb_plan_routes_host", "mpcb_plan_last_error", "mpcb_step_static", "mpcb_session_step_all_host",
]

F_STATE, F_OA, F_OD, F_TRAJ, F_XREF, F_DI_AI, F_XDEV, F_TARGET, F_WARM, F_STATUS, F_ITERS, F_DONE, F_NCOURSE, \
    F_ROUTE, F_GOAL, F_AGENT_IDX, F_COLLIDED, F_COLL_IDX, F_SCRIPT, F_OBS, F_NOBS, F_LAUNCH_ORDER = range(22)
MAX_OBSTACLES = 8
STATIC_MAX_BATCH = 16384


def build(verbose=False):
    """Compile libmpcb.so in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    srcs = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith(".cu")]
    deps = srcs + [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cuh")] + \
        [os.path.join(_HERE, "..", "include", "mpcb.h")]
    if os.path.exists(LIB_PATH) and all(os.path.getmtime(LIB_PATH) >= os.path.getmtime(d) for d in deps):
        return LIB_PATH
    extra = os.environ.get("MPCB_EXTRA_NVCC_FLAGS", "").split()
    cmd = ["nvcc"] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB_PATH] + srcs
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise MpcbError("nvcc failed:\n" + res.stderr[-4000:])
    if verbose:
        print(res.stderr)
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise MpcbError(f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                        "(there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp, ip, dp = C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_double)
    L.mpcb_version.restype = C.c_int
    L.mpcb_default_config.argtypes = [C.POINTER(mpcb_config)]
    L.mpcb_default_config.restype = None
    L.mpcb_device_count.restype = C.c_int
    L.mpcb_create.argtypes = [C.POINTER(mpcb_config), C.c_int, C.c_int, C.POINTER(vp)]
    L.mpcb_create.restype = C.c_int
    L.mpcb_destroy.argtypes = [vp]
    L.mpcb_destroy.restype = None
    L.mpcb_last_error.argtypes = [vp]
    L.mpcb_last_error.restype = C.c_char_p
    L.mpcb_supported_horizons.argtypes = [ip, C.c_int]
    L.mpcb_supported_horizons.restype = C.c_int
    L.mpcb_smooth_yaw.argtypes = [vp, C.c_int]
    L.mpcb_smooth_yaw.restype = None
    L.mpcb_set_courses_host.argtypes = [vp, vp, vp, vp, C.c_int, C.c_int]
    L.mpcb_set_courses_host.restype = C.c_int
    L.mpcb_step.argtypes = [vp, C.c_int] + [vp] * 15
    L.mpcb_step.restype = C.c_int
    L.mpcb_step_static.argtypes = [vp, C.c_int] + [vp] * 15
    L.mpcb_step_static.restype = C.c_int
    L.mpcb_step_host.argtypes = [vp, C.c_int] + [vp] * 14
    L.mpcb_step_host.restype = C.c_int
    L.mpcb_plant_step.argtypes = [vp, C.c_int, vp, vp, vp, vp, vp]
    L.mpcb_plant_step.restype = C.c_int
    L.mpcb_is_goal.argtypes = [vp, C.c_int, vp, vp, vp, vp, vp, vp]
    L.mpcb_is_goal.restype = C.c_int
    L.mpcb_session_begin.argtypes = [vp, C.c_int, vp, vp, vp, vp]
    L.mpcb_session_begin.restype = C.c_int
    L.mpcb_session_set_ncourse_host.argtypes = [vp, vp]
    L.mpcb_session_set_ncourse_host.restype = C.c_int
    L.mpcb_session_step_host.argtypes = [vp, vp, vp, vp]
    L.mpcb_session_step_host.restype = C.c_int
    L.mpcb_session_step_all_host.argtypes = [vp] * 12
    L.mpcb_session_step_all_host.restype = C.c_int
    L.mpcb_session_step_dev.argtypes = [vp, vp]
    L.mpcb_session_step_dev.restype = C.c_int
    L.mpcb_session_closed_loop_dev.argtypes = [vp, C.c_int, vp, C.c_int, vp]
    L.mpcb_session_closed_loop_dev.restype = C.c_int
    L.mpcb_session_device_ptr.argtypes = [vp, C.c_int]
    L.mpcb_session_device_ptr.restype = vp
    L.mpcb_session_field_bytes.argtypes = [vp, C.c_int]
    L.mpcb_session_field_bytes.restype = C.c_longlong
    L.mpcb_session_fetch_host.argtypes = [vp, C.c_int, vp, C.c_longlong]
    L.mpcb_session_fetch_host.restype = C.c_int
    L.mpcb_session_store_host.argtypes = [vp, C.c_int, vp, C.c_longlong]
    L.mpcb_session_store_host.restype = C.c_int
    L.mpcb_coupling.argtypes = [vp, C.c_int] + [vp] * 12 + [C.c_int, vp]
    L.mpcb_coupling.restype = C.c_int
    L.mpcb_coupling_host.argtypes = [vp, C.c_int] + [vp] * 11 + [C.c_int]
    L.mpcb_coupling_host.restype = C.c_int
    L.mpcb_session_multi_agent_dev.argtypes = [vp, C.c_int, C.c_int, vp, C.c_int, vp]
    L.mpcb_session_multi_agent_dev.restype = C.c_int
    L.mpcb_session_set_scripted_host.argtypes = [vp, vp, vp]
    L.mpcb_session_set_scripted_host.restype = C.c_int
    L.mpcb_session_scripted_loop_dev.argtypes = [vp, C.c_int, vp, vp, C.c_int, vp]
    L.mpcb_session_scripted_loop_dev.restype = C.c_int
    L.mpcb_session_set_weights_host.argtypes = [vp, vp]
    L.mpcb_session_set_weights_host.restype = C.c_int
    L.mpcb_session_set_speed_cutoff_host.argtypes = [vp, vp]
    L.mpcb_session_set_speed_cutoff_host.restype = C.c_int
    L.mpcb_plan_routes_host.argtypes = [C.c_int, C.c_int, vp, vp, vp, vp, vp, C.c_int, vp, vp, vp, C.c_int, C.c_int, vp, vp, vp,
                                        C.c_int, vp, C.c_int, C.c_int, C.c_int] + [vp] * 9
    L.mpcb_plan_routes_host.restype = C.c_int
    L.mpcb_plan_last_error.argtypes = []
    L.mpcb_plan_last_error.restype = C.c_char_p
    L.mpcb_launch_count.argtypes = [vp]
    L.mpcb_launch_count.restype = C.c_longlong
    L.mpcb_measure_fp64_peak.argtypes = [C.c_int, dp]
    L.mpcb_measure_fp64_peak.restype = C.c_int
    _lib = L
    return L


def check(rc, handle=None, what=""):
    if rc != 0:
        msg = lib().mpcb_last_error(handle)
        raise MpcbError(f"{what} failed (rc={rc}): {msg.decode() if msg else ''}")
