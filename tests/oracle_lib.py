"""ctypes access to the test oracle (oracle/liboracle.so) and, when present, to the
reference build (oracle/_ref/libransac_ref.so). TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")


class Options(C.Structure):
    _fields_ = [("min_iters", C.c_uint32), ("max_iters", C.c_uint32), ("success_prob", C.c_double),
                ("sq_thr", C.c_double), ("seed", C.c_uint32), ("lo_steps", C.c_int32),
                ("thr_mult", C.c_double), ("lsq_iters", C.c_int32), ("min_sample_mult", C.c_int32),
                ("non_min_sample_mult", C.c_int32), ("lo_start", C.c_uint32), ("final_lsq", C.c_int32)]

    @staticmethod
    def default(**kw):
        o = Options(100, 10000, 0.9999, 1.0, 0, 10, 2.0 ** 0.5, 4, 7, 3, 50, 0)
        for k, v in kw.items():
            setattr(o, k, v)
        return o


class Stats(C.Structure):
    _fields_ = [("num_iterations", C.c_uint32), ("best_num_inliers", C.c_int32),
                ("best_model_score", C.c_double), ("inlier_ratio", C.c_double),
                ("number_lo_iterations", C.c_int32), ("found", C.c_int32)]


class HybridOptions(C.Structure):
    _fields_ = [("min_iters", C.c_uint32), ("max_iters", C.c_uint32), ("max_iters_per_solver", C.c_uint32),
                ("success_prob", C.c_double), ("sq_thr", C.c_double * 2), ("weights", C.c_double * 2),
                ("seed", C.c_uint32), ("lo_steps", C.c_int32), ("thr_mult", C.c_double),
                ("lsq_iters", C.c_int32), ("min_sample_mult", C.c_int32), ("lo_start", C.c_uint32),
                ("final_lsq", C.c_int32)]

    @staticmethod
    def default(**kw):
        o = HybridOptions(100, 10000, 10000, 0.9999, (C.c_double * 2)(1.0, 1.0), (C.c_double * 2)(1.0, 1.0),
                          0, 10, 2.0 ** 0.5, 4, 7, 50, 0)
        for k, v in kw.items():
            if k in ("sq_thr", "weights"):
                v = (C.c_double * 2)(*v)
            setattr(o, k, v)
        return o


class HybridStats(C.Structure):
    _fields_ = [("num_iterations_total", C.c_uint32), ("num_iterations_per_solver", C.c_uint32 * 2),
                ("best_num_inliers", C.c_int32), ("best_solver_type", C.c_int32),
                ("best_model_score", C.c_double), ("inlier_ratios", C.c_double * 2),
                ("num_inliers_per_type", C.c_int32 * 2), ("number_lo_iterations", C.c_int32)]


def _bind(lib, prefix):
    def fn(name, res, args):
        f = getattr(lib, prefix + name)
        f.restype = res
        f.argtypes = args
        return f

    ns = type("OracleFns", (), {})()
    ns.sample_uniform = fn("sample_uniform", None, [C.c_uint32, C.c_int, C.c_int, C.c_int, i32p])
    ns.sample_hybrid = fn("sample_hybrid", None, [C.c_uint32, C.c_int, i32p, C.c_int, i32p, C.c_int, i32p])
    ns.sample_hybrid_biased = fn("sample_hybrid_biased", None,
                                 [C.c_uint32, C.c_int, i32p, f64p, C.c_int, i32p, C.c_int, i32p])
    ns.num_required_iterations = fn("num_required_iterations", C.c_uint32,
                                    [C.c_double, C.c_double, C.c_int, C.c_uint32, C.c_uint32])
    ns.num_required_iterations_hybrid = fn("num_required_iterations_hybrid", C.c_uint32,
                                           [C.c_int, f64p, C.c_double, i32p, C.c_uint32, C.c_uint32])
    ns.shuffle_and_resize = fn("shuffle_and_resize", None,
                               [C.c_uint32, C.c_int, C.c_int, i32p, C.POINTER(C.c_uint32)])
    ns.shuffle_and_resize_reps = fn("shuffle_and_resize_reps", None,
                                    [C.c_uint32, C.c_int, C.c_int, C.c_int, i32p, C.POINTER(C.c_uint32)])
    ns.pixels_to_rays = fn("pixels_to_rays", None, [C.c_double, C.c_double, C.c_int, f64p, f64p])
    ns.pose_errors = fn("pose_errors", None, [f64p, C.c_double, C.c_double, C.c_int, f64p, f64p, f64p])
    ns.pose_score_models = fn("pose_score_models", None,
                              [C.c_double, C.c_double, C.c_int, f64p, f64p, C.c_int, f64p, C.c_double, f64p, i32p])
    ns.pose_inliers = fn("pose_inliers", C.c_int,
                         [C.c_double, C.c_double, C.c_int, f64p, f64p, f64p, C.c_double, i32p])
    ns.p3p = fn("p3p", C.c_int, [f64p, f64p, f64p, f64p])
    ns.pose_minimal_batch = fn("pose_minimal_batch", None,
                               [C.c_double, C.c_double, C.c_double, C.c_int, f64p, f64p, C.c_int, i32p, f64p, u8p])
    ns.pose_least_squares = fn("pose_least_squares", None,
                               [C.c_double, C.c_double, C.c_int, f64p, f64p, C.c_int, i32p, f64p])
    ns.lomsac_pose = fn("lomsac_pose", None,
                        [C.POINTER(Options), C.c_double, C.c_double, C.c_int, f64p, f64p, f64p,
                         C.POINTER(Stats), i32p])
    ns.time_pose_scoring = fn("time_pose_scoring", C.c_double,
                              [C.c_double, C.c_double, C.c_int, f64p, f64p, C.c_int, f64p, C.c_double, C.c_int,
                               C.POINTER(C.c_double)])
    ns.line_score_models = fn("line_score_models", None, [C.c_int, f64p, C.c_int, f64p, C.c_double, f64p, i32p])
    ns.line_inliers = fn("line_inliers", C.c_int, [C.c_int, f64p, f64p, C.c_double, i32p])
    ns.line_minimal_batch = fn("line_minimal_batch", None, [C.c_int, f64p, C.c_int, i32p, f64p, u8p])
    ns.line_least_squares = fn("line_least_squares", None, [C.c_int, f64p, C.c_int, i32p, f64p, C.POINTER(C.c_int32)])
    ns.lomsac_line = fn("lomsac_line", None, [C.POINTER(Options), C.c_int, f64p, f64p, C.POINTER(Stats), i32p])
    ns.lomsac_circle = fn("lomsac_circle", None, [C.POINTER(Options), C.c_int, f64p, f64p, C.POINTER(Stats), i32p])
    ns.hybrid_line_score_models = fn("hybrid_line_score_models", None,
                                     [C.c_int, f64p, C.c_int, f64p, C.c_int, f64p, f64p, f64p, f64p, i32p])
    ns.hybrid_line_inliers = fn("hybrid_line_inliers", C.c_int,
                                [C.c_int, f64p, C.c_int, f64p, f64p, f64p, i32p, i32p, i32p])
    ns.point_normal_batch = fn("point_normal_batch", None, [C.c_int, f64p, C.c_int, i32p, f64p, u8p])
    ns.hybrid_lomsac_line = fn("hybrid_lomsac_line", None,
                               [C.POINTER(HybridOptions), C.c_int, f64p, C.c_int, f64p, f64p, f64p,
                                C.POINTER(HybridStats), i32p, i32p])
    ns.hybrid_lomsac_line_biased = fn("hybrid_lomsac_line_biased", None,
                                      [C.POINTER(HybridOptions), C.c_int, f64p, f64p, C.c_int, f64p, f64p, f64p, f64p,
                                       C.POINTER(HybridStats), i32p, i32p])
    ns.hybrid_lomsac_pose = fn("hybrid_lomsac_pose", None,
                               [C.POINTER(HybridOptions), C.c_double, C.c_double, C.c_int, f64p, f64p, C.c_int, f64p,
                                f64p, f64p, f64p, C.POINTER(HybridStats), i32p, i32p])
    ns.hybrid_pose_errors = fn("hybrid_pose_errors", None,
                               [f64p, C.c_double, C.c_double, C.c_int, f64p, f64p, C.c_int, f64p, f64p, f64p, f64p])
    return ns


def build_oracle():
    subprocess.run(["make", "-s", "-C", ORACLE_DIR], check=True, capture_output=True)


_cache = {}


def oracle():
    """The restated oracle (always available: built from committed sources)."""
    if "o" not in _cache:
        path = os.path.join(ORACLE_DIR, "liboracle.so")
        if not os.path.exists(path):
            build_oracle()
        _cache["o"] = _bind(C.CDLL(path), "oracle_")
    return _cache["o"]


def _host_has_avx512():
    try:
        with open("/proc/cpuinfo") as fh:
            for ln in fh:
                if ln.startswith("flags"):
                    fl = set(ln.split(":", 1)[1].split())
                    return {"avx512f", "avx512dq", "avx512cd", "avx512bw", "avx512vl"} <= fl
    except OSError:
        pass
    return False


def reference_fast_flags():
    """Which -march the timed reference build in use was compiled for (the widest this host supports: what the
    reference's own -march=native, examples/CMakeLists.txt:7,15, selects on this host)."""
    v4 = os.path.join(ORACLE_DIR, "_ref", "libransac_ref_fast_v4.so")
    return "x86-64-v4" if (_host_has_avx512() and os.path.exists(v4)) else "x86-64-v3"


def reference(fast=False):
    """The reference's own headers compiled in oracle/_ref, or None if not built."""
    key = "rf" if fast else "r"
    if key not in _cache:
        name = "libransac_ref.so"
        if fast:
            name = "libransac_ref_fast_v4.so" if reference_fast_flags() == "x86-64-v4" else "libransac_ref_fast.so"
        path = os.path.join(ORACLE_DIR, "_ref", name)
        _cache[key] = _bind(C.CDLL(path), "ref_") if os.path.exists(path) else None
    return _cache[key]


# ---- thin numpy helpers used by many tests ------------------------------------
def c64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def pose_score(lib, fx, fy, p2d, p3d, models, thr):
    models = c64(models).reshape(-1, 12)
    k = len(models)
    scores = np.empty(k)
    counts = np.empty(k, np.int32)
    lib.pose_score_models(fx, fy, len(p2d), c64(p2d), c64(p3d), k, models, thr, scores, counts)
    return scores, counts


def pose_inliers(lib, fx, fy, p2d, p3d, model, thr):
    idx = np.empty(len(p2d), np.int32)
    c = lib.pose_inliers(fx, fy, len(p2d), c64(p2d), c64(p3d), c64(model), thr, idx)
    return idx[:c].copy()


def pose_minimal(lib, fx, fy, thr, p2d, p3d, samples):
    samples = np.ascontiguousarray(samples, np.int32).reshape(-1, 4)
    k = len(samples)
    models = np.empty((k, 12))
    valid = np.empty(k, np.uint8)
    lib.pose_minimal_batch(fx, fy, thr, len(p2d), c64(p2d), c64(p3d), k, samples, models, valid)
    return models, valid


def sample_uniform(lib, seed, n, s, k):
    out = np.empty((k, s), np.int32)
    lib.sample_uniform(seed, n, s, k, out)
    return out


def hybrid_lomsac_pose(lib, opts, fx, fy, p2d, p3d, cam1, world1, priors):
    """Hybrid LO-MSAC over 2D-3D matches (type 0) and 3D-3D matches (type 1) with oracle/solvers_oracle.h:
    HybridPoseSolver. Returns (model (12,), HybridStats, inliers type 0, inliers type 1)."""
    n0, n1 = len(p2d), len(cam1)
    m = np.zeros(12)
    st = HybridStats()
    a, b = np.zeros(max(n0, 1), np.int32), np.zeros(max(n1, 1), np.int32)
    lib.hybrid_lomsac_pose(C.byref(opts), fx, fy, n0, c64(p2d), c64(p3d), n1, c64(cam1), c64(world1),
                           np.asarray(priors, dtype=np.float64), m, C.byref(st), a, b)
    return m, st, a[:st.num_inliers_per_type[0]].copy(), b[:st.num_inliers_per_type[1]].copy()
