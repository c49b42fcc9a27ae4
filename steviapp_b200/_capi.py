"""ctypes binding of libsba_b200.so (include/sba_b200.h).

The library is the product; this module only declares its C ABI.  It fails
loudly when the shared object is missing -- there is no Python or CPU fallback.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libsba_b200.so")

SBA_MAX_HISTORY = 512
SBA_OK = 0
SBA_ERR_NO_DEVICE = -1
SBA_ERR_CUDA = -2
SBA_ERR_INVALID = -3
SBA_ERR_EVAL_FAILED = -4
SBA_ERR_NUMERIC = -5
SBA_ERR_COMM = -6

EXPORTS = [
    "sba_default_options", "sba_create", "sba_destroy", "sba_last_error", "sba_create_error",
    "sba_set_intrinsics", "sba_set_poses", "sba_set_points", "sba_add_observations", "sba_add_priors",
    "sba_finalize", "sba_upload_parameters", "sba_download_parameters", "sba_evaluate",
    "sba_evaluate_intrinsics_jacobian", "sba_evaluate_offset", "sba_normal_blocks", "sba_schur_matvec", "sba_solve", "sba_set_iteration_callback", "sba_set_options",
    "sba_time_stage", "sba_triangulate", "sba_comm_unique_id", "sba_comm_init", "sba_partition_points",
    "sba_problem_sizes", "sba_explicit_info", "sba_explicit_shard", "sba_explicit_sparse_info", "sba_explicit_tiles", "sba_version",
    "sba_bal_open", "sba_bal_close", "sba_bal_error", "sba_bal_sizes", "sba_bal_raw", "sba_bal_convert", "sba_bal_write", "sba_bal_cost", "sba_bal_cameras_from_engine",
    "sba_pushbroom_equivalent_uv",
    "sba_set_intrinsics_blocks", "sba_add_intrinsics_priors", "sba_set_sensor_offsets", "sba_set_lever_arms", "sba_shared_info",
    "sba_point_covariances", "sba_add_uv_pairs", "sba_set_pair_loss", "sba_evaluate_pairs",
    "sba_project_open", "sba_project_close", "sba_project_error", "sba_project_sizes", "sba_project_ids", "sba_project_convert",
    "sba_project_write_results", "sba_project_save",
    "sba_triangulate_noalloc",
]


class SbaOptions(C.Structure):
    _fields_ = [
        ("device", C.c_int32), ("max_iterations", C.c_int32),
        ("function_tolerance", C.c_double), ("parameter_tolerance", C.c_double), ("gradient_tolerance", C.c_double),
        ("initial_radius", C.c_double), ("max_radius", C.c_double), ("min_radius", C.c_double),
        ("min_lm_diagonal", C.c_double), ("max_lm_diagonal", C.c_double), ("min_relative_decrease", C.c_double),
        ("jacobi_scaling", C.c_int32), ("pcg_rel_tol", C.c_double), ("pcg_max_iters", C.c_int32),
        ("pcg_check_every", C.c_int32), ("operator_mode", C.c_int32), ("verbose", C.c_int32),
        ("linear_solver", C.c_int32), ("pcg_eta", C.c_double),
    ]


class SbaSummary(C.Structure):
    _fields_ = [
        ("status", C.c_int32), ("iterations", C.c_int32), ("successful_steps", C.c_int32),
        ("pcg_iterations_total", C.c_int32),
        ("initial_cost", C.c_double), ("final_cost", C.c_double),
        ("cost_history", C.c_double * SBA_MAX_HISTORY), ("radius_history", C.c_double * SBA_MAX_HISTORY),
        ("pcg_iters_history", C.c_int32 * SBA_MAX_HISTORY),
        ("ms_total", C.c_double), ("ms_linearize", C.c_double), ("ms_schur", C.c_double), ("ms_pcg", C.c_double),
        ("ms_update", C.c_double), ("ms_h2d", C.c_double), ("ms_d2h", C.c_double),
        ("kernel_launches", C.c_int32),
    ]


class SbaIterationInfo(C.Structure):
    _fields_ = [
        ("iteration", C.c_int32), ("cost", C.c_double), ("cost_change", C.c_double), ("radius", C.c_double),
        ("step_valid", C.c_int32), ("step_accepted", C.c_int32), ("pcg_iterations", C.c_int32),
    ]


ITERATION_CALLBACK = C.CFUNCTYPE(C.c_int, C.POINTER(SbaIterationInfo), C.c_void_p)

_lib = None


def load():
    """dlopen libsba_b200.so; raises if it was not built (run build.sh / __graft_entry__.build())."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with ./build.sh (nvcc, sm_100a). "
            "steviapp_b200 has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    vp, i32, dbl = C.c_void_p, C.c_int32, C.c_double
    P = C.POINTER
    lib.sba_default_options.argtypes = [P(SbaOptions)]
    lib.sba_default_options.restype = None
    lib.sba_create.argtypes = [P(SbaOptions), P(vp)]
    lib.sba_destroy.argtypes = [vp]
    lib.sba_destroy.restype = None
    lib.sba_last_error.argtypes = [vp]
    lib.sba_last_error.restype = C.c_char_p
    lib.sba_create_error.restype = C.c_char_p
    lib.sba_version.restype = C.c_char_p
    lib.sba_set_options.argtypes = [vp, P(SbaOptions)]
    lib.sba_set_intrinsics.argtypes = [vp, i32, vp]
    lib.sba_set_poses.argtypes = [vp, vp, vp, i32, vp, vp, vp]
    lib.sba_set_points.argtypes = [vp, vp, i32, vp]
    lib.sba_add_observations.argtypes = [vp, i32, vp, vp, vp, vp]
    lib.sba_add_priors.argtypes = [vp, i32, vp, vp, vp, vp]
    lib.sba_finalize.argtypes = [vp]
    lib.sba_upload_parameters.argtypes = [vp]
    lib.sba_download_parameters.argtypes = [vp]
    lib.sba_evaluate.argtypes = [vp, P(dbl), vp, vp, vp]
    lib.sba_normal_blocks.argtypes = [vp, vp, vp, vp, vp]
    lib.sba_schur_matvec.argtypes = [vp, vp, vp, vp, vp, vp]
    lib.sba_solve.argtypes = [vp, i32, P(SbaSummary)]
    lib.sba_set_iteration_callback.argtypes = [vp, ITERATION_CALLBACK, vp]
    lib.sba_time_stage.argtypes = [vp, i32, i32, P(dbl)]
    lib.sba_triangulate.argtypes = [vp, i32, vp, vp, vp, vp, vp, P(dbl)]
    lib.sba_comm_unique_id.argtypes = [vp]
    lib.sba_comm_init.argtypes = [vp, i32, i32, vp]
    lib.sba_partition_points.argtypes = [i32, vp, i32, vp]
    lib.sba_problem_sizes.argtypes = [vp, P(i32), P(i32), P(i32), P(i32), P(i32)]
    lib.sba_explicit_info.argtypes = [vp, P(i32), P(i32), P(i32), P(i32)]
    lib.sba_evaluate_intrinsics_jacobian.argtypes = [vp, vp]
    lib.sba_explicit_shard.argtypes = [vp, P(i32), P(i32), P(i32), P(i32)]
    lib.sba_explicit_sparse_info.argtypes = [vp, P(i32), P(i32), P(i32), P(C.c_int64), P(C.c_int64)]
    lib.sba_explicit_tiles.argtypes = [vp, vp, vp, vp, vp]
    lib.sba_evaluate_offset.argtypes = [vp, vp, vp, vp, vp, vp]
    lib.sba_set_intrinsics_blocks.argtypes = [vp, i32, vp, vp]
    lib.sba_add_intrinsics_priors.argtypes = [vp, i32, vp, vp, vp, vp]
    lib.sba_set_sensor_offsets.argtypes = [vp, vp]
    lib.sba_set_lever_arms.argtypes = [vp, i32, vp, vp, vp, vp]
    lib.sba_shared_info.argtypes = [vp, P(i32), P(i32), vp, vp]
    lib.sba_add_uv_pairs.argtypes = [vp, i32, vp, vp, vp, vp, vp, vp]
    lib.sba_set_pair_loss.argtypes = [vp, i32, dbl]
    lib.sba_evaluate_pairs.argtypes = [vp, vp, vp]
    lib.sba_point_covariances.argtypes = [vp, vp]
    lib.sba_bal_open.argtypes = [C.c_char_p, C.POINTER(vp)]
    lib.sba_bal_close.argtypes = [vp]
    lib.sba_bal_close.restype = None
    lib.sba_bal_error.restype = C.c_char_p
    lib.sba_bal_sizes.argtypes = [vp, vp, vp, vp]
    lib.sba_bal_raw.argtypes = [vp, vp, vp, vp, vp, vp]
    lib.sba_bal_convert.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]
    lib.sba_bal_cost.argtypes = [vp, vp, vp, P(dbl), P(dbl)]
    lib.sba_bal_cameras_from_engine.argtypes = [vp, vp, vp, vp]
    lib.sba_pushbroom_equivalent_uv.argtypes = [C.c_int32, vp, C.c_double, C.c_double, C.c_double, vp, vp, vp, vp]
    lib.sba_bal_write.argtypes = [C.c_char_p, C.c_int32, C.c_int32, C.c_int32, vp, vp, vp, vp, vp]
    lib.sba_project_open.argtypes = [C.c_char_p, C.POINTER(vp)]
    lib.sba_project_close.argtypes = [vp]
    lib.sba_project_close.restype = None
    lib.sba_project_error.restype = C.c_char_p
    lib.sba_project_sizes.argtypes = [vp] + [P(i32)] * 7
    lib.sba_project_ids.argtypes = [vp, vp, vp, vp]
    lib.sba_project_convert.argtypes = [vp, i32] + [vp] * 20 + [P(i32)]
    lib.sba_project_write_results.argtypes = [vp, vp, vp, vp, vp]
    lib.sba_project_save.argtypes = [vp, C.c_char_p]
    lib.sba_triangulate_noalloc.argtypes = [i32, vp, vp, vp, i32, vp, i32, i32, vp, vp, vp, vp, vp, vp]
    for name in EXPORTS:
        fn = getattr(lib, name)
        if fn.restype is C.c_int:   # default restype: all status-returning functions
            fn.restype = C.c_int
    _lib = lib
    return lib
