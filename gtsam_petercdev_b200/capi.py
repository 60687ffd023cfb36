"""ctypes binding of include/gtsam_b200.h.  Loads the in-tree libgtsam_b200.so and fails loudly if it is missing:
there is no Python or CPU fallback for any numeric entry point."""
import ctypes as C
import os
import re

HERE = os.path.dirname(os.path.abspath(__file__))
# GTB_LIB: another build of the same library (A/B timing of two commits on one box); default: the in-tree build
LIB_PATH = os.environ.get("GTB_LIB") or os.path.join(HERE, "libgtsam_b200.so")
HEADER = os.path.join(os.path.dirname(HERE), "include", "gtsam_b200.h")

OK, ERR_INVALID, ERR_NO_DEVICE, ERR_CUDA, ERR_INDETERMINATE, ERR_KEY_MISSING, ERR_NOISE_MODEL, ERR_INCONSISTENT = range(8)


class LMParams(C.Structure):
    _fields_ = [("max_iterations", C.c_int32), ("relative_error_tol", C.c_double), ("absolute_error_tol", C.c_double),
                ("error_tol", C.c_double), ("lambda_initial", C.c_double), ("lambda_factor", C.c_double),
                ("lambda_upper_bound", C.c_double), ("lambda_lower_bound", C.c_double),
                ("min_model_fidelity", C.c_double), ("diagonal_damping", C.c_int32),
                ("use_fixed_lambda_factor", C.c_int32), ("min_diagonal", C.c_double), ("max_diagonal", C.c_double)]


class LMResult(C.Structure):
    _fields_ = [("initial_error", C.c_double), ("final_error", C.c_double), ("final_lambda", C.c_double),
                ("iterations", C.c_int32), ("inner_iterations", C.c_int32), ("log_rows", C.c_int32)]


class ISAM2Result(C.Structure):
    _fields_ = [("update_count", C.c_int64), ("variables_relinearized", C.c_int64), ("n_variables", C.c_int64),
                ("n_factors", C.c_int64), ("n_cliques", C.c_int64), ("plan_ms", C.c_double), ("solve_ms", C.c_double)]


_P = C.c_void_p
_SIGS = {
    "gtb_create": (C.c_int, [C.c_int, C.POINTER(_P)]),
    "gtb_destroy": (None, [_P]),
    "gtb_last_error": (C.c_char_p, [_P]),
    "gtb_version": (C.c_char_p, []),
    "gtb_set_variables": (C.c_int, [_P, C.c_int, C.c_int64, _P, _P]),
    "gtb_set_factors": (C.c_int, [_P, C.c_int, C.c_int64, _P, _P, _P, _P]),
    "gtb_set_graph_order": (C.c_int, [_P, C.c_int64, _P]),
    "gtb_finalize_problem": (C.c_int, [_P]),
    "gtb_set_ordering": (C.c_int, [_P, C.c_int64, _P]),
    "gtb_set_ordering_builtin": (C.c_int, [_P, C.c_int]),
    "gtb_plan_sizes": (C.c_int, [_P, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "gtb_plan_get": (C.c_int, [_P, _P, _P, _P, _P]),
    "gtb_plan_stats": (C.c_int, [_P, _P]),
    "gtb_linearize": (C.c_int, [_P]),
    "gtb_get_jacobians": (C.c_int, [_P, C.c_int, _P]),
    "gtb_hessian_diagonal": (C.c_int, [_P, _P]),
    "gtb_factorize": (C.c_int, [_P, C.c_double, C.c_int, C.c_double, C.c_double, C.POINTER(C.c_uint64)]),
    "gtb_backsolve": (C.c_int, [_P]),
    "gtb_trial": (C.c_int, [_P, C.c_double, C.c_int, C.c_double, C.c_double, _P, C.POINTER(C.c_uint64)]),
    "gtb_get_delta": (C.c_int, [_P, _P]),
    "gtb_get_conditionals": (C.c_int, [_P, _P]),
    "gtb_linear_errors": (C.c_int, [_P, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "gtb_retract": (C.c_int, [_P]),
    "gtb_error": (C.c_int, [_P, C.c_int, C.POINTER(C.c_double)]),
    "gtb_accept": (C.c_int, [_P]),
    "gtb_get_variables": (C.c_int, [_P, C.c_int, C.c_int, _P]),
    "gtb_lm_params_default": (None, [C.POINTER(LMParams)]),
    "gtb_lm_params_ceres": (None, [C.POINTER(LMParams)]),
    "gtb_lm_optimize": (C.c_int, [_P, C.POINTER(LMParams), C.POINTER(LMResult), _P, C.c_int64]),
    "gtb_lm_iterate": (C.c_int, [_P, C.POINTER(LMParams), C.POINTER(C.c_double), C.POINTER(C.c_double),
                                 C.POINTER(C.c_double), C.POINTER(C.c_int32)]),
    "gtb_lm_last_accepted": (C.c_int, [_P]),
    "gtb_gn_optimize": (C.c_int, [_P, C.c_int32, C.c_double, C.c_double, C.POINTER(LMResult)]),
    "gtb_kernel_launches": (C.c_int64, [_P]),
    "gtb_stage_ms": (C.c_int, [_P, _P]),
    "gtb_set_stage_timing": (C.c_int, [_P, C.c_int]),
    "gtb_profile": (C.c_int, [_P, C.c_int]),
    "gtb_profile_classes": (C.c_int, []),
    "gtb_profile_class_name": (C.c_char_p, [C.c_int]),
    "gtb_profile_read": (C.c_int, [_P, _P, _P]),
    "gtb_profile_read_raw": (C.c_int64, [_P, _P, _P, C.c_int64]),
    "gtb_stream": (_P, [_P]),
    "gtb_set_robust": (C.c_int, [_P, C.c_int, C.c_int, C.c_double]),
    "gtb_set_distributed": (C.c_int, [_P, C.c_int32, C.c_int32, _P, _P]),
    "gtb_partition_get": (C.c_int, [_P, _P, _P]),
    "gtb_schedule_check": (C.c_int, [_P, _P]),
    "gtb_marginals_compute": (C.c_int, [_P]),
    "gtb_marginal_covariance": (C.c_int, [_P, C.c_int64, _P, _P]),
    "gtb_dataset_read_g2o": (C.c_int, [C.c_char_p, C.c_int, C.c_int, C.POINTER(_P)]),
    "gtb_dataset_read_bal": (C.c_int, [C.c_char_p, C.POINTER(_P)]),
    "gtb_dataset_free": (None, [_P]),
    "gtb_dataset_last_error": (C.c_char_p, []),
    "gtb_dataset_num_variables": (C.c_int64, [_P, C.c_int]),
    "gtb_dataset_variable_keys": (_P, [_P, C.c_int]),
    "gtb_dataset_variable_state": (_P, [_P, C.c_int]),
    "gtb_dataset_num_factors": (C.c_int64, [_P, C.c_int]),
    "gtb_dataset_factor_keys": (_P, [_P, C.c_int]),
    "gtb_dataset_factor_z": (_P, [_P, C.c_int]),
    "gtb_dataset_factor_noise_kind": (_P, [_P, C.c_int]),
    "gtb_dataset_factor_noise": (_P, [_P, C.c_int]),
    "gtb_dataset_robust": (C.c_int, [_P, C.POINTER(C.c_double)]),
    "gtb_dataset_apply": (C.c_int, [_P, _P]),
    "gtb_relinearize_marked": (C.c_int, [_P, C.c_double, _P, C.POINTER(C.c_int64)]),
    "gtb_isam2_create": (C.c_int, [C.c_int, C.c_double, C.c_int32, C.POINTER(_P)]),
    "gtb_isam2_destroy": (None, [_P]),
    "gtb_isam2_last_error": (C.c_char_p, [_P]),
    "gtb_isam2_add_variables": (C.c_int, [_P, C.c_int, C.c_int64, _P, _P]),
    "gtb_isam2_add_factors": (C.c_int, [_P, C.c_int, C.c_int64, _P, _P, _P, _P]),
    "gtb_isam2_set_ordering": (C.c_int, [_P, C.c_int64, _P]),
    "gtb_isam2_update": (C.c_int, [_P, C.POINTER(ISAM2Result), C.POINTER(C.c_uint64)]),
    "gtb_isam2_estimate": (C.c_int, [_P, C.c_int, _P]),
    "gtb_isam2_count": (C.c_int64, [_P, C.c_int]),
    "gtb_isam2_engine": (_P, [_P]),
    "gtb_noalloc_setup": (C.c_int, [C.c_int]),
    "gtb_bucket_error_noalloc": (C.c_int, [C.c_int, C.c_int, C.c_int64, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "gtb_noalloc_last_error": (C.c_char_p, []),
    "gtb_write_g2o": (C.c_int, [C.c_char_p, C.c_int64, _P, _P, C.c_int64, _P, _P, C.c_int64, _P, _P, _P, C.c_int64, _P, _P, _P]),
}

# int (*gtb_comm_fn)(void* user, int op, void* buf, int64_t count, void* stream)
COMM_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int64, C.c_void_p)
COMM_REDUCE_SUM_TO_0, COMM_BROADCAST_FROM_0, COMM_ALLREDUCE_SUM, COMM_ALLREDUCE_MIN = 0, 1, 2, 3

_lib = None


def header_symbols():
    """Every function name declared in include/gtsam_b200.h."""
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(gtb_[a-z0-9_]+)\s*\(", text)))


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: build it with `python -m gtsam_petercdev_b200.build` "
                              "(there is no fallback implementation)")
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib
