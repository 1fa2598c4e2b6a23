"""ctypes binding of include/ransaclib_b200.h (the C ABI of the CUDA library).

The shared library is built in-tree by ``__graft_entry__.build()`` /
``make -C ransaclib_b200/csrc``. There is no fallback: if the library is missing or
no CUDA device is present, loading / context creation raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libransaclib_b200.so")

RLB_F64, RLB_F32, RLB_F64_PLAIN, RLB_F64_FILTERED, RLB_F32_SORTED, RLB_F32_UNSORTED = 0, 1, 2, 3, 4, 5
KIND_LINE, KIND_HYBRID_LINE, KIND_POSE = 0, 1, 2

# every symbol include/ransaclib_b200.h declares (tests check the .so exports all of them)
SYMBOLS = [
    "rlb_last_error", "rlb_version", "rlb_ctx_create", "rlb_ctx_destroy", "rlb_ctx_set_stream", "rlb_ctx_use_own_stream",
    "rlb_ctx_synchronize", "rlb_ctx_launch_count", "rlb_ctx_sm_count", "rlb_scorer_create_line",
    "rlb_scorer_create_hybrid_line", "rlb_scorer_create_pose", "rlb_scorer_create_custom", "rlb_scorer_create_custom_hybrid",
    "rlb_custom_compile_check", "rlb_scorer_destroy", "rlb_scorer_kind",
    "rlb_scorer_num_types", "rlb_scorer_num_data", "rlb_scorer_model_size", "rlb_scorer_get_rays",
    "rlb_score_models", "rlb_score_models_dev", "rlb_get_inliers", "rlb_get_inliers_view", "rlb_get_inliers_dev",
    "rlb_eval_errors", "rlb_solve_line2pt_batch", "rlb_solve_point_normal_batch", "rlb_solve_p3p_batch",
    "rlb_hypothesize_and_score", "rlb_hypothesize_and_score_dev", "rlb_measure_fma_peak", "rlb_sharded_round_dev", "rlb_select_winner_dev", "rlb_ctx_set_timing", "rlb_ctx_kernel_time_ms", "rlb_ctx_last_sweep_plan",
    "rlb_localize_default_options", "rlb_localize_batch", "rlb_localize_batch_dev",
    "rlb_lomsac_default_options", "rlb_line_lomsac_batch", "rlb_line_lomsac_batch_dev", "rlb_scorer_line_lomsac",
    "rlb_scorer_create_pose_dev", "rlb_get_inliers_noalloc_dev",
    "rlb_comm_get_unique_id", "rlb_comm_init_rank", "rlb_comm_create_local", "rlb_comm_destroy", "rlb_comm_nranks",
    "rlb_comm_num_local", "rlb_comm_rank", "rlb_comm_ctx", "rlb_comm_collective_count",
    "rlb_comm_scorer_create_pose", "rlb_comm_scorer_destroy", "rlb_comm_scorer_local",
    "rlb_comm_hypothesize_and_score", "rlb_comm_hypothesize_and_score_models", "rlb_comm_round_dev", "rlb_comm_get_inliers",
]
COMM_ID_BYTES = 128
NOALLOC_MAX_POINTS = 4194304


class LocalizeOptions(C.Structure):
    """rlb_localize_options (include/ransaclib_b200.h)."""
    _fields_ = [("min_iters", C.c_uint32), ("max_iters", C.c_uint32), ("success_prob", C.c_double),
                ("sq_thr", C.c_double), ("lo_steps", C.c_int32), ("thr_mult", C.c_double),
                ("lsq_iters", C.c_int32), ("min_sample_mult", C.c_int32), ("non_min_sample_mult", C.c_int32),
                ("lo_start", C.c_uint32), ("final_lsq", C.c_int32), ("debug_flags", C.c_int32),
                ("handback_after", C.c_uint32), ("handback_min_remaining", C.c_uint32)]


# rlb_localize_result as a numpy record (40 bytes)
LOCALIZE_RESULT = np.dtype([("success", np.int32), ("num_inliers", np.int32), ("num_iterations", np.uint32),
                            ("num_lo", np.int32), ("score", np.float64), ("inlier_ratio", np.float64),
                            ("passes", np.int32), ("lsq_calls", np.int32)], align=True)


class RlbError(RuntimeError):
    pass


_lib = None


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RlbError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C ransaclib_b200/csrc`. ransaclib_b200 has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    vp, i32, f64 = C.c_void_p, C.c_int, C.c_double
    lib.rlb_last_error.restype = C.c_char_p
    lib.rlb_version.restype = i32
    lib.rlb_ctx_create.argtypes = [i32, C.POINTER(vp)]
    lib.rlb_ctx_destroy.argtypes = [vp]
    lib.rlb_ctx_set_stream.argtypes = [vp, vp]
    lib.rlb_ctx_use_own_stream.argtypes = [vp]
    lib.rlb_ctx_synchronize.argtypes = [vp]
    lib.rlb_ctx_launch_count.argtypes = [vp]
    lib.rlb_ctx_launch_count.restype = C.c_int64
    lib.rlb_ctx_sm_count.argtypes = [vp]
    lib.rlb_scorer_create_line.argtypes = [vp, i32, vp, C.POINTER(vp)]
    lib.rlb_scorer_create_hybrid_line.argtypes = [vp, i32, vp, i32, vp, C.POINTER(vp)]
    lib.rlb_scorer_create_pose.argtypes = [vp, i32, vp, vp, f64, f64, C.POINTER(vp)]
    lib.rlb_scorer_destroy.argtypes = [vp]
    for n in ("rlb_scorer_kind", "rlb_scorer_num_types", "rlb_scorer_model_size"):
        getattr(lib, n).argtypes = [vp]
    lib.rlb_scorer_num_data.argtypes = [vp, i32]
    lib.rlb_scorer_get_rays.argtypes = [vp, vp]
    lib.rlb_score_models.argtypes = [vp, i32, i32, vp, vp, vp, vp, vp, vp, vp, vp]
    lib.rlb_score_models_dev.argtypes = [vp, i32, i32, vp, vp, vp, vp, vp, vp, vp]
    lib.rlb_get_inliers.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp]
    lib.rlb_get_inliers_view.argtypes = [vp, vp, vp, vp, vp, C.POINTER(vp), vp, vp]
    lib.rlb_get_inliers_dev.argtypes = [vp, vp, vp, vp, vp, vp, vp]
    lib.rlb_eval_errors.argtypes = [vp, i32, vp, i32, vp]
    lib.rlb_solve_line2pt_batch.argtypes = [vp, i32, vp, vp, vp]
    lib.rlb_solve_point_normal_batch.argtypes = [vp, i32, vp, vp, vp]
    lib.rlb_solve_p3p_batch.argtypes = [vp, i32, vp, f64, vp, vp]
    lib.rlb_hypothesize_and_score.argtypes = [vp, i32, i32, i32, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]
    lib.rlb_hypothesize_and_score_dev.argtypes = [vp, i32, i32, i32, vp, vp, vp, vp, vp, vp, vp, vp]
    lib.rlb_measure_fma_peak.argtypes = [vp, i32, C.POINTER(f64)]
    lib.rlb_sharded_round_dev.argtypes = [vp, i32, i32, i32, vp, vp, vp, C.c_int64, vp, vp, vp, vp, vp]
    lib.rlb_select_winner_dev.argtypes = [vp, vp, i32, i32, vp]
    lib.rlb_ctx_set_timing.argtypes = [vp, i32]
    lib.rlb_ctx_kernel_time_ms.argtypes = [vp, i32, C.POINTER(f64)]
    lib.rlb_ctx_last_sweep_plan.argtypes = [vp, C.POINTER(i32), C.POINTER(i32)]
    lib.rlb_scorer_create_custom.argtypes = [vp, i32, i32, vp, i32, C.c_char_p, C.POINTER(vp)]
    lib.rlb_scorer_create_custom_hybrid.argtypes = [vp, i32, i32, vp, i32, i32, vp, i32, C.c_char_p, C.POINTER(vp)]
    lib.rlb_custom_compile_check.argtypes = [C.c_char_p, i32, i32, C.POINTER(C.c_size_t)]
    lib.rlb_localize_default_options.argtypes = [C.POINTER(LocalizeOptions), f64, i32, C.c_uint32, C.c_uint32]
    lib.rlb_localize_default_options.restype = None
    lib.rlb_localize_batch.argtypes = [vp, i32, vp, vp, vp, vp, vp, vp, C.POINTER(LocalizeOptions), vp, vp, vp]
    lib.rlb_localize_batch_dev.argtypes = [vp, i32, i32, vp, vp, vp, vp, vp, vp, C.POINTER(LocalizeOptions), vp,
                                           vp, vp]
    lib.rlb_lomsac_default_options.argtypes = [C.POINTER(LocalizeOptions)]
    lib.rlb_lomsac_default_options.restype = None
    lib.rlb_line_lomsac_batch.argtypes = [vp, i32, vp, vp, vp, C.POINTER(LocalizeOptions), vp, vp, vp]
    lib.rlb_line_lomsac_batch_dev.argtypes = [vp, i32, i32, vp, vp, vp, C.POINTER(LocalizeOptions), vp, vp, vp]
    lib.rlb_scorer_line_lomsac.argtypes = [vp, C.POINTER(LocalizeOptions), C.c_uint32, vp, vp, vp]
    lib.rlb_scorer_create_pose_dev.argtypes = [vp, i32, vp, vp, f64, f64, C.POINTER(vp)]
    lib.rlb_get_inliers_noalloc_dev.argtypes = [vp, vp, vp, vp, vp, vp, vp]
    lib.rlb_comm_get_unique_id.argtypes = [vp]
    lib.rlb_comm_init_rank.argtypes = [vp, i32, i32, vp, C.POINTER(vp)]
    lib.rlb_comm_create_local.argtypes = [i32, vp, C.POINTER(vp)]
    lib.rlb_comm_destroy.argtypes = [vp]
    lib.rlb_comm_nranks.argtypes = [vp]
    lib.rlb_comm_num_local.argtypes = [vp]
    lib.rlb_comm_rank.argtypes = [vp, i32]
    lib.rlb_comm_ctx.argtypes = [vp, i32]
    lib.rlb_comm_ctx.restype = vp
    lib.rlb_comm_collective_count.argtypes = [vp]
    lib.rlb_comm_collective_count.restype = C.c_int64
    lib.rlb_comm_scorer_create_pose.argtypes = [vp, i32, vp, vp, f64, f64, i32, C.POINTER(vp)]
    lib.rlb_comm_scorer_destroy.argtypes = [vp]
    lib.rlb_comm_scorer_local.argtypes = [vp, i32]
    lib.rlb_comm_scorer_local.restype = vp
    lib.rlb_comm_hypothesize_and_score.argtypes = [vp, i32, i32, i32, vp, vp, vp, vp, vp, vp, vp, vp]
    lib.rlb_comm_round_dev.argtypes = [vp, i32, i32, i32, vp, vp, vp, C.c_int64, vp, vp, vp, vp, vp]
    lib.rlb_comm_get_inliers.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp]
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        raise RlbError(f"ransaclib_b200 error {rc}: {load().rlb_last_error().decode()}")


def ptr(a):
    """Host pointer of a numpy array (or None)."""
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def f64a(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a if shape is None else a.reshape(shape)


def i32a(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.int32)
    return a if shape is None else a.reshape(shape)
