"""ctypes binding of libm3r_b200.so (the C ABI declared in include/m3r_backends.h).

There is no CPU fallback: if the CUDA library is missing the import fails loudly.
Build it with `python __graft_entry__.py build` (or `python lightweight-mast3r-slam_b200/build.py`).
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get(
    "M3R_B200_LIB", os.path.join(os.path.dirname(_HERE), "libm3r_b200.so")
)

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"mast3r_slam_backends: CUDA library {LIB_PATH} not found. There is no CPU fallback; "
        "build it with `python __graft_entry__.py build`."
    )

lib = ctypes.CDLL(LIB_PATH)

c_int, c_float, c_void_p, c_int64 = ctypes.c_int, ctypes.c_float, ctypes.c_void_p, ctypes.c_int64

M3R_OK = 0
DTYPE_F16, DTYPE_F32, DTYPE_F64 = 0, 1, 2
GN_RAYS, GN_CALIB, GN_POINTS = 0, 1, 2


class GnProblem(ctypes.Structure):
    """Mirror of `m3r_gn_problem` (include/m3r_backends.h)."""

    _fields_ = [
        ("kind", c_int),
        ("num_poses", c_int),
        ("num_points", c_int),
        ("num_edges", c_int),
        ("edge_begin", c_int),
        ("edge_end", c_int),
        ("Twc", c_void_p),
        ("Xs", c_void_p),
        ("Cs", c_void_p),
        ("ii", c_void_p),
        ("jj", c_void_p),
        ("idx_ii2jj", c_void_p),
        ("valid_match", c_void_p),
        ("Q", c_void_p),
        ("sigma_ray", c_float),
        ("sigma_dist", c_float),
        ("K", c_void_p),
        ("height", c_int),
        ("width", c_int),
        ("pixel_border", c_int),
        ("z_eps", c_float),
        ("sigma_pixel", c_float),
        ("sigma_depth", c_float),
        ("sigma_point", c_float),
        ("C_thresh", c_float),
        ("Q_thresh", c_float),
    ]


class TrackParams(ctypes.Structure):
    """Mirror of `m3r_track_params` (include/m3r_backends.h)."""

    _fields_ = [
        ("kind", c_int),
        ("max_iters", c_int),
        ("rel_error", c_float),
        ("delta_norm", c_float),
        ("huber_k", c_float),
        ("sigma_a", c_float),
        ("sigma_b", c_float),
        ("height", c_int),
        ("width", c_int),
        ("pixel_border", c_float),
        ("z_eps", c_float),
    ]


class TrackResult(ctypes.Structure):
    """Mirror of `m3r_track_result` (include/m3r_backends.h); lives in device memory."""

    _fields_ = [
        ("T_WCf", c_float * 8),
        ("T_CkCf", c_float * 8),
        ("status", ctypes.c_int32),
        ("iterations", ctypes.c_int32),
        ("cost", ctypes.c_double),
    ]


TRACK_OK, TRACK_CHOLESKY_FAILED = 0, 1


def _sig(name, restype, argtypes):
    fn = getattr(lib, name)
    fn.restype = restype
    fn.argtypes = argtypes
    return fn


abi_version = _sig("m3r_abi_version", c_int, [])
last_error = _sig("m3r_last_error", ctypes.c_char_p, [])
launch_count = _sig("m3r_launch_count", ctypes.c_uint64, [])

iter_proj = _sig(
    "m3r_iter_proj",
    c_int,
    [c_void_p] * 5 + [c_int] * 5 + [c_float, c_float, c_void_p],
)
iter_proj_v = _sig(
    "m3r_iter_proj_v",
    c_int,
    [c_void_p] * 5 + [c_int] * 5 + [c_float, c_float, c_int, c_void_p],
)
refine_matches = _sig(
    "m3r_refine_matches", c_int, [c_void_p] * 4 + [c_int] * 8 + [c_void_p]
)
refine_matches_v = _sig(
    "m3r_refine_matches_v", c_int, [c_void_p] * 4 + [c_int] * 9 + [c_void_p]
)
refine_matches_static = _sig(
    "m3r_refine_matches_static", c_int, [c_void_p] * 4 + [c_int] * 8 + [c_void_p]
)
refine_matches_static_max_tiles = _sig("m3r_refine_matches_static_max_tiles", c_int64, [])
prep_rays = _sig("m3r_prep_rays", c_int, [c_void_p] * 2 + [c_int] * 3 + [c_void_p])
match = _sig(
    "m3r_match",
    c_int,
    [c_void_p] * 4 + [c_int] * 2 + [c_void_p] * 5 + [c_int] * 4 + [c_float] * 3 + [c_int] * 2 + [c_void_p],
)
match_confidence_scratch_bytes = _sig("m3r_match_confidence_scratch_bytes", c_int64, [c_int, c_int])
match_confidence = _sig(
    "m3r_match_confidence",
    c_int,
    [c_void_p] * 6 + [c_float] * 2 + [c_void_p] * 5 + [c_int] * 3 + [c_void_p],
)
track_scratch_doubles = _sig("m3r_track_scratch_doubles", c_int64, [])
track_pose = _sig(
    "m3r_track_pose",
    c_int,
    [ctypes.POINTER(TrackParams), c_void_p, c_int64] + [c_void_p] * 7 + [c_int64] + [c_void_p] * 5,
)
update_pointmap = _sig("m3r_update_pointmap", c_int, [c_int] + [c_void_p] * 5 + [c_int64, c_void_p])
constrain_points_to_ray = _sig(
    "m3r_constrain_points_to_ray", c_int, [c_void_p] * 3 + [c_int] * 3 + [c_void_p]
)
FILTER_MODES = {"weighted_pointmap": 0, "weighted_spherical": 1, "indep_conf": 2}

gn_workspace_create = _sig("m3r_gn_workspace_create", c_int, [ctypes.POINTER(c_void_p)])
gn_workspace_destroy = _sig("m3r_gn_workspace_destroy", None, [c_void_p])

gauss_newton_rays = _sig(
    "m3r_gauss_newton_rays",
    c_int,
    [c_void_p] * 10 + [c_int] * 3 + [c_float] * 4 + [c_int, c_float, c_void_p],
)
gauss_newton_calib = _sig(
    "m3r_gauss_newton_calib",
    c_int,
    [c_void_p] * 11 + [c_int] * 6 + [c_float] * 5 + [c_int, c_float, c_void_p],
)
gauss_newton_points = _sig(
    "m3r_gauss_newton_points",
    c_int,
    [c_void_p] * 10 + [c_int] * 3 + [c_float] * 3 + [c_int, c_float, c_void_p],
)

gn_system_size = _sig("m3r_gn_system_size", c_int64, [c_int])
gn_begin = _sig("m3r_gn_begin", c_int, [c_void_p, ctypes.POINTER(GnProblem), c_void_p])
gn_linearize = _sig(
    "m3r_gn_linearize", c_int, [c_void_p, ctypes.POINTER(GnProblem), c_void_p, c_void_p]
)
gn_solve_retract = _sig(
    "m3r_gn_solve_retract",
    c_int,
    [c_void_p, ctypes.POINTER(GnProblem), c_void_p, c_void_p, c_float, c_void_p],
)
gn_edge_blocks = _sig(
    "m3r_gn_edge_blocks", c_int, [c_void_p, ctypes.POINTER(c_void_p), ctypes.POINTER(c_void_p)]
)
gn_iterations_done = _sig(
    "m3r_gn_iterations_done", c_int, [c_void_p, c_void_p, ctypes.POINTER(c_int)]
)

IPC_HANDLE_BYTES = 64
gn_comm_create = _sig("m3r_gn_comm_create", c_int, [ctypes.POINTER(c_void_p), c_int, c_int, c_int64])
gn_comm_destroy = _sig("m3r_gn_comm_destroy", None, [c_void_p])
gn_comm_handle = _sig("m3r_gn_comm_handle", c_int, [c_void_p, c_void_p])
gn_comm_connect = _sig("m3r_gn_comm_connect", c_int, [c_void_p, c_void_p])
gn_comm_connect_local = _sig("m3r_gn_comm_connect_local", c_int, [ctypes.POINTER(c_void_p), c_int])
gn_comm_next_round = _sig("m3r_gn_comm_next_round", c_int, [c_void_p, ctypes.POINTER(c_void_p)])
gn_comm_signal = _sig("m3r_gn_comm_signal", c_int, [c_void_p, c_void_p, c_void_p])
gn_comm_reduce = _sig("m3r_gn_comm_reduce", c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p])
gn_comm_reduce_phase = _sig(
    "m3r_gn_comm_reduce_phase", c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p]
)
debug_set_exchange_mode = _sig("m3r_debug_set_exchange_mode", None, [c_int])
gn_multi_create = _sig("m3r_gn_multi_create", c_int, [ctypes.POINTER(c_void_p), ctypes.POINTER(c_int), c_int])
gn_multi_destroy = _sig("m3r_gn_multi_destroy", None, [c_void_p])
gn_multi_solve = _sig(
    "m3r_gn_multi_solve", c_int, [c_void_p, ctypes.POINTER(GnProblem), c_void_p, c_int, c_float, c_void_p]
)
gn_multi_report = _sig(
    "m3r_gn_multi_report",
    c_int,
    [c_void_p, c_void_p, c_int, ctypes.POINTER(c_float), ctypes.POINTER(c_int), ctypes.POINTER(c_int)],
)

# every symbol include/m3r_backends.h declares (checked by tests/test_abi.py)
EXPORTS = [
    "m3r_abi_version",
    "m3r_last_error",
    "m3r_launch_count",
    "m3r_iter_proj",
    "m3r_iter_proj_v",
    "m3r_refine_matches",
    "m3r_refine_matches_v",
    "m3r_refine_matches_static",
    "m3r_refine_matches_static_max_tiles",
    "m3r_prep_rays",
    "m3r_match",
    "m3r_match_confidence_scratch_bytes",
    "m3r_match_confidence",
    "m3r_track_scratch_doubles",
    "m3r_track_pose",
    "m3r_update_pointmap",
    "m3r_constrain_points_to_ray",
    "m3r_gn_workspace_create",
    "m3r_gn_workspace_destroy",
    "m3r_gauss_newton_rays",
    "m3r_gauss_newton_calib",
    "m3r_gauss_newton_points",
    "m3r_gn_system_size",
    "m3r_gn_begin",
    "m3r_gn_linearize",
    "m3r_gn_solve_retract",
    "m3r_gn_edge_blocks",
    "m3r_gn_iterations_done",
    "m3r_gn_comm_create",
    "m3r_gn_comm_destroy",
    "m3r_gn_comm_handle",
    "m3r_gn_comm_connect",
    "m3r_gn_comm_connect_local",
    "m3r_gn_comm_next_round",
    "m3r_gn_comm_signal",
    "m3r_gn_comm_reduce",
    "m3r_gn_comm_reduce_phase",
    "m3r_debug_set_exchange_mode",
    "m3r_gn_multi_create",
    "m3r_gn_multi_destroy",
    "m3r_gn_multi_solve",
    "m3r_gn_multi_report",
]


def check(rc, what):
    if rc != M3R_OK:
        msg = last_error()
        raise RuntimeError(f"{what}: {msg.decode() if msg else 'error'} (code {rc})")
