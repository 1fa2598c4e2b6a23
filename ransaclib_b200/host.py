"""Python mirror of the drop-in template layer: LO-MSAC / Hybrid LO-MSAC ``EstimateModel``
for the three example solvers, executed by ``libransaclib_b200_host.so`` (include/RansacLib/*.h
over the CUDA C ABI). Same option names as the reference's ``LORansacOptions`` /
``HybridLORansacOptions`` (RansacLib/ransac.h:47-92, hybrid_ransac.h:47-93)."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import _capi

_HOST_PATH = os.environ.get("RLB_HOST_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib",
                                                           "libransaclib_b200_host.so")


class _Options(C.Structure):
    _fields_ = [("min_iters", C.c_uint32), ("max_iters", C.c_uint32), ("success_prob", C.c_double),
                ("sq_thr", C.c_double), ("seed", C.c_uint32), ("lo_steps", C.c_int32),
                ("thr_mult", C.c_double), ("lsq_iters", C.c_int32), ("min_sample_mult", C.c_int32),
                ("non_min_sample_mult", C.c_int32), ("lo_start", C.c_uint32), ("final_lsq", C.c_int32)]


class _Stats(C.Structure):
    _fields_ = [("num_iterations", C.c_uint32), ("best_num_inliers", C.c_int32),
                ("best_model_score", C.c_double), ("inlier_ratio", C.c_double),
                ("number_lo_iterations", C.c_int32), ("found", C.c_int32), ("batches", C.c_uint64),
                ("hypotheses_scored", C.c_uint64), ("single_passes", C.c_uint64), ("seconds", C.c_double),
                ("handbacks", C.c_uint64), ("handbacks_resumed", C.c_uint64)]


class _HybridOptions(C.Structure):
    _fields_ = [("min_iters", C.c_uint32), ("max_iters", C.c_uint32), ("max_iters_per_solver", C.c_uint32),
                ("success_prob", C.c_double), ("sq_thr", C.c_double * 2), ("weights", C.c_double * 2),
                ("seed", C.c_uint32), ("lo_steps", C.c_int32), ("thr_mult", C.c_double),
                ("lsq_iters", C.c_int32), ("min_sample_mult", C.c_int32), ("lo_start", C.c_uint32),
                ("final_lsq", C.c_int32)]


class _HybridStats(C.Structure):
    _fields_ = [("num_iterations_total", C.c_uint32), ("num_iterations_per_solver", C.c_uint32 * 2),
                ("best_num_inliers", C.c_int32), ("best_solver_type", C.c_int32),
                ("best_model_score", C.c_double), ("inlier_ratios", C.c_double * 2),
                ("num_inliers_per_type", C.c_int32 * 2), ("number_lo_iterations", C.c_int32)]


_lib = None


def _load():
    global _lib
    if _lib is None:
        _capi.load()  # the CUDA library first (the host library links against it)
        if not os.path.exists(_HOST_PATH):
            raise _capi.RlbError(f"{_HOST_PATH} is missing: run __graft_entry__.build()")
        _lib = C.CDLL(_HOST_PATH)
        _lib.rlbh_last_error.restype = C.c_char_p
    return _lib


def _check(rc):
    if rc < 0:
        raise _capi.RlbError(_load().rlbh_last_error().decode())


class LORansacOptions:
    """Defaults of reference ransac.h:47-92."""

    def __init__(self, **kw):
        self.min_num_iterations_ = 100
        self.max_num_iterations_ = 10000
        self.success_probability_ = 0.9999
        self.squared_inlier_threshold_ = 1.0
        self.random_seed_ = 0
        self.num_lo_steps_ = 10
        self.threshold_multiplier_ = 2.0 ** 0.5
        self.num_lsq_iterations_ = 4
        self.min_sample_multiplicator_ = 7
        self.non_min_sample_multiplier_ = 3
        self.lo_starting_iterations_ = 50
        self.final_least_squares_ = False
        for k, v in kw.items():
            if not hasattr(self, k):
                raise AttributeError(k)
            setattr(self, k, v)

    def _c(self):
        return _Options(self.min_num_iterations_, self.max_num_iterations_, self.success_probability_,
                        self.squared_inlier_threshold_, self.random_seed_, self.num_lo_steps_,
                        self.threshold_multiplier_, self.num_lsq_iterations_, self.min_sample_multiplicator_,
                        self.non_min_sample_multiplier_, self.lo_starting_iterations_,
                        int(self.final_least_squares_))


class HybridLORansacOptions:
    """Defaults of reference hybrid_ransac.h:47-93."""

    def __init__(self, **kw):
        self.min_num_iterations_ = 100
        self.max_num_iterations_ = 10000
        self.max_num_iterations_per_solver_ = 10000
        self.success_probability_ = 0.9999
        self.squared_inlier_thresholds_ = [1.0, 1.0]
        self.data_type_weights_ = [1.0, 1.0]
        self.random_seed_ = 0
        self.num_lo_steps_ = 10
        self.threshold_multiplier_ = 2.0 ** 0.5
        self.num_lsq_iterations_ = 4
        self.min_sample_multiplicator_ = 7
        self.lo_starting_iterations_ = 50
        self.final_least_squares_ = False
        for k, v in kw.items():
            if not hasattr(self, k):
                raise AttributeError(k)
            setattr(self, k, v)

    def _c(self):
        return _HybridOptions(self.min_num_iterations_, self.max_num_iterations_,
                              self.max_num_iterations_per_solver_, self.success_probability_,
                              (C.c_double * 2)(*self.squared_inlier_thresholds_),
                              (C.c_double * 2)(*self.data_type_weights_), self.random_seed_, self.num_lo_steps_,
                              self.threshold_multiplier_, self.num_lsq_iterations_,
                              self.min_sample_multiplicator_, self.lo_starting_iterations_,
                              int(self.final_least_squares_))


def _stats_dict(st, inliers):
    return {"num_iterations": st.num_iterations, "best_num_inliers": st.best_num_inliers,
            "best_model_score": st.best_model_score, "inlier_ratio": st.inlier_ratio,
            "number_lo_iterations": st.number_lo_iterations, "found": bool(st.found),
            "inlier_indices": inliers[:st.best_num_inliers].copy(), "batches": st.batches,
            "hypotheses_scored": st.hypotheses_scored, "single_passes": st.single_passes,
            "seconds": st.seconds, "handbacks": st.handbacks, "handbacks_resumed": st.handbacks_resumed}


def estimate_line(options: LORansacOptions, xy, device=0, precision=-1):
    """LocallyOptimizedMSAC<Line, ., LineEstimator>::EstimateModel -> (model[3], stats)."""
    xy = _capi.f64a(xy, (-1, 2))
    o, st = options._c(), _Stats()
    model = np.zeros(3)
    inl = np.zeros(max(len(xy), 1), np.int32)
    _check(_load().rlbh_lomsac_line(C.byref(o), device, precision, len(xy), _capi.ptr(xy), _capi.ptr(model),
                                    C.byref(st), _capi.ptr(inl)))
    return model, _stats_dict(st, inl)


def estimate_circle(options: LORansacOptions, xy, device=0):
    """LocallyOptimizedMSAC<Circle2D, ., CircleEstimator>::EstimateModel -> ((cx, cy, r), stats): a solver with a
    user-defined residual (examples/circle_estimator.h), host minimal solver, f64 sweeps."""
    xy = _capi.f64a(xy, (-1, 2))
    o, st = options._c(), _Stats()
    model = np.zeros(3)
    inl = np.zeros(max(len(xy), 1), np.int32)
    _check(_load().rlbh_lomsac_circle(C.byref(o), device, len(xy), _capi.ptr(xy), _capi.ptr(model), C.byref(st),
                                      _capi.ptr(inl)))
    return model, _stats_dict(st, inl)


def estimate_pose(options: LORansacOptions, fx, fy, points2D, points3D, device=0, precision=-1):
    """LocallyOptimizedMSAC<CameraPose, ., CalibratedAbsolutePoseEstimator>::EstimateModel."""
    p2 = _capi.f64a(points2D, (-1, 2))
    p3 = _capi.f64a(points3D, (-1, 3))
    o, st = options._c(), _Stats()
    model = np.zeros(12)
    inl = np.zeros(max(len(p2), 1), np.int32)
    _check(_load().rlbh_lomsac_pose(C.byref(o), device, precision, C.c_double(fx), C.c_double(fy), len(p2),
                                    _capi.ptr(p2), _capi.ptr(p3), _capi.ptr(model), C.byref(st), _capi.ptr(inl)))
    return model, _stats_dict(st, inl)


def pose_minimal_solver(fx, fy, sq_thr, points2D, points3D, samples, device=0):
    """CalibratedAbsolutePoseEstimator::MinimalSolver (the host-side P3P of examples/p3p.h) on k samples.
    -> (models k x 12, valid k); an invalid sample's model is zero, like rlb_solve_p3p_batch."""
    p2 = _capi.f64a(points2D, (-1, 2))
    p3 = _capi.f64a(points3D, (-1, 3))
    smp = _capi.i32a(samples, (-1, 4))
    models = np.zeros((len(smp), 12))
    valid = np.zeros(len(smp), np.uint8)
    _check(_load().rlbh_pose_minimal_solver(device, C.c_double(fx), C.c_double(fy), C.c_double(sq_thr), len(p2),
                                            _capi.ptr(p2), _capi.ptr(p3), len(smp), _capi.ptr(smp),
                                            _capi.ptr(models), _capi.ptr(valid)))
    return models, valid


def estimate_hybrid_line(options: HybridLORansacOptions, xy0, xynn1, priors, device=0, precision=-1,
                         sampling_weights=None, custom_functor=False):
    """HybridLocallyOptimizedMSAC<Line, ., HybridLineEstimator[, Sampler]>::EstimateModel. sampling_weights =
    (w0, w1) per-point weights selects HybridBiasedSampling (reference hybrid_sampling.h:146-223) as the Sampler."""
    xy0 = _capi.f64a(xy0, (-1, 2))
    p1 = _capi.f64a(xynn1, (-1, 4))
    pr = _capi.f64a(priors, (2,))
    o, st = options._c(), _HybridStats()
    model = np.zeros(3)
    a = np.zeros(max(len(xy0), 1), np.int32)
    b = np.zeros(max(len(p1), 1), np.int32)
    if custom_functor:  # examples/hybrid_custom_line_estimator.h: the residual reaches the GPU as source
        _check(_load().rlbh_hybrid_lomsac_line_custom(C.byref(o), device, precision, len(xy0), _capi.ptr(xy0),
                                                      len(p1), _capi.ptr(p1), _capi.ptr(pr), _capi.ptr(model),
                                                      C.byref(st), _capi.ptr(a), _capi.ptr(b)))
    elif sampling_weights is None:
        _check(_load().rlbh_hybrid_lomsac_line(C.byref(o), device, precision, len(xy0), _capi.ptr(xy0), len(p1),
                                               _capi.ptr(p1), _capi.ptr(pr), _capi.ptr(model), C.byref(st),
                                               _capi.ptr(a), _capi.ptr(b)))
    else:
        w0 = _capi.f64a(sampling_weights[0], (len(xy0),))
        w1 = _capi.f64a(sampling_weights[1], (len(p1),))
        _check(_load().rlbh_hybrid_lomsac_line_biased(C.byref(o), device, precision, len(xy0), _capi.ptr(xy0),
                                                      _capi.ptr(w0), len(p1), _capi.ptr(p1), _capi.ptr(w1),
                                                      _capi.ptr(pr), _capi.ptr(model), C.byref(st), _capi.ptr(a),
                                                      _capi.ptr(b)))
    return model, {
        "num_iterations_total": st.num_iterations_total,
        "num_iterations_per_solver": list(st.num_iterations_per_solver),
        "best_num_inliers": st.best_num_inliers, "best_solver_type": st.best_solver_type,
        "best_model_score": st.best_model_score, "inlier_ratios": list(st.inlier_ratios),
        "number_lo_iterations": st.number_lo_iterations,
        "inlier_indices": [a[:st.num_inliers_per_type[0]].copy(), b[:st.num_inliers_per_type[1]].copy()]}


def estimate_hybrid_pose(options: HybridLORansacOptions, fx, fy, points2D, points3D, cam1, world1, priors, device=0,
                         precision=-1):
    """HybridLocallyOptimizedMSAC<CameraPose, ., HybridPoseEstimator>::EstimateModel (examples/hybrid_pose_estimator.h):
    2D-3D matches (data type 0) and 3D-3D matches (data type 1: a point measured in the camera frame against its
    world position), P3P and three-point minimal solvers; both residuals reach the GPU as source (a user-defined
    HybridSolver). Returns (model: R column-major then c, stats)."""
    p2 = _capi.f64a(points2D, (-1, 2))
    p3 = _capi.f64a(points3D, (-1, 3))
    c1 = _capi.f64a(cam1, (-1, 3))
    w1 = _capi.f64a(world1, (-1, 3))
    if len(p2) != len(p3) or len(c1) != len(w1):
        raise ValueError("matches come in pairs")
    pr = _capi.f64a(priors, (2,))
    o, st = options._c(), _HybridStats()
    model = np.zeros(12)
    a = np.zeros(max(len(p2), 1), np.int32)
    b = np.zeros(max(len(c1), 1), np.int32)
    _check(_load().rlbh_hybrid_lomsac_pose(C.byref(o), device, precision, C.c_double(fx), C.c_double(fy), len(p2),
                                           _capi.ptr(p2), _capi.ptr(p3), len(c1), _capi.ptr(c1), _capi.ptr(w1),
                                           _capi.ptr(pr), _capi.ptr(model), C.byref(st), _capi.ptr(a), _capi.ptr(b)))
    return model, {
        "num_iterations_total": st.num_iterations_total,
        "num_iterations_per_solver": list(st.num_iterations_per_solver),
        "best_num_inliers": st.best_num_inliers, "best_solver_type": st.best_solver_type,
        "best_model_score": st.best_model_score, "inlier_ratios": list(st.inlier_ratios),
        "number_lo_iterations": st.number_lo_iterations,
        "inlier_indices": [a[:st.num_inliers_per_type[0]].copy(), b[:st.num_inliers_per_type[1]].copy()]}


def sample_uniform(seed, n, s, k):
    """k draws of the template layer's UniformSampling (include/RansacLib/sampling.h); host only."""
    out = np.empty((k, s), np.int32)
    _check(_load().rlbh_sample_uniform(C.c_uint32(seed), n, s, k, _capi.ptr(out)))
    return out


def shuffle_and_resize(seed, n, target, reps=1):
    """`reps` consecutive RandomShuffleAndResize(target, ., iota(n)) on the drivers' own engine (include/RansacLib/
    utils.h: Mt19937, block-wise consumption of the draws that cannot reach the kept prefix); host only.
    -> (kept prefix of the last shuffle, next raw engine output, first 8 engine outputs after seeding)."""
    out = np.empty(max(target, 1), np.int32)
    nxt = C.c_uint32()
    first = np.empty(8, np.uint32)
    _check(_load().rlbh_shuffle_and_resize(C.c_uint32(seed), n, target, reps, _capi.ptr(out), C.byref(nxt),
                                           _capi.ptr(first)))
    return out[:target].copy(), nxt.value, first


def shuffle_from_list(seed, n, target, reps=1):
    """The same through utils::ShuffleSampleFromList on the read-only list 3 i + 1 (the drivers' shuffle over
    the pinned result list of a fused pass). -> (sample of the last shuffle, next raw engine output)."""
    out = np.empty(max(target, 1), np.int32)
    nxt = C.c_uint32()
    _check(_load().rlbh_shuffle_from_list(C.c_uint32(seed), n, target, reps, _capi.ptr(out), C.byref(nxt)))
    return out[:target].copy(), nxt.value


def sample_hybrid(seed, num_data, sizes, weights=None):
    """Draws of HybridUniformSampling (weights None) / HybridBiasedSampling (per-type weight lists) from
    include/RansacLib/hybrid_sampling.h; sizes is k x T. Returns a list of k lists of T index lists."""
    nd = np.ascontiguousarray(num_data, np.int32)
    sz = np.ascontiguousarray(sizes, np.int32).reshape(-1, len(nd))
    mx = max(int(sz.max()), 1)
    out = np.empty((len(sz), len(nd), mx), np.int32)
    w = None
    if weights is not None:
        w = np.ascontiguousarray(np.concatenate([np.asarray(x, np.float64) for x in weights]))
    _check(_load().rlbh_sample_hybrid(C.c_uint32(seed), len(nd), _capi.ptr(nd), _capi.ptr(w) if w is not None else None,
                                      len(sz), _capi.ptr(sz), mx, _capi.ptr(out)))
    return [[[int(v) for v in row if v >= 0] for row in draw] for draw in out]


def localize(focal_x, focal_y, points2D, points3D, inlier_threshold, number_lo_steps, min_num_iterations,
             max_num_iterations, random_seed=0, device=0, precision=-1):
    """Core of pyransaclib.ransaclib_localization -> dict (see ransaclib_b200/pyransaclib.py)."""
    p2 = _capi.f64a(points2D, (-1, 2))
    p3 = _capi.f64a(points3D, (-1, 3))
    n = len(p2)
    q, t = np.zeros(4), np.zeros(3)
    ninl = C.c_int32()
    inl = np.zeros(max(n, 1), np.int32)
    st = _Stats()
    rc = _load().rlbh_localize(device, precision, C.c_float(focal_x), C.c_float(focal_y), n, _capi.ptr(p2),
                               _capi.ptr(p3), C.c_float(inlier_threshold), int(number_lo_steps),
                               int(min_num_iterations), int(max_num_iterations), int(random_seed), _capi.ptr(q),
                               _capi.ptr(t), C.byref(ninl), _capi.ptr(inl), C.byref(st))
    _check(rc)
    if rc == 0:
        return {"success": False}
    return {"success": True, "qvec": q, "tvec": t, "num_inliers": int(ninl.value),
            "inliers": inl[:ninl.value].tolist(), "_stats": _stats_dict(st, inl)}


def localize_batch(focal_x, focal_y, points2D_list, points3D_list, inlier_threshold, number_lo_steps,
                   min_num_iterations, max_num_iterations, random_seeds, device=0):
    """Q independent queries through ONE launch of the device-resident LO-MSAC
    (rlbh_localize_batch -> rlb_localize_batch). -> list of dicts like ``localize``."""
    nq = len(points2D_list)
    p2s = [_capi.f64a(a, (-1, 2)) for a in points2D_list]
    p3s = [_capi.f64a(a, (-1, 3)) for a in points3D_list]
    off = np.zeros(nq + 1, np.int64)
    for q in range(nq):
        if len(p2s[q]) != len(p3s[q]):
            raise ValueError(f"query {q}: points2D and points3D differ in length")
        off[q + 1] = off[q] + len(p2s[q])
    total = int(off[-1])
    p2 = np.concatenate(p2s) if total else np.zeros((0, 2))
    p3 = np.concatenate(p3s) if total else np.zeros((0, 3))
    fx = np.ascontiguousarray(np.broadcast_to(np.asarray(focal_x, np.float32), (nq,)))
    fy = np.ascontiguousarray(np.broadcast_to(np.asarray(focal_y, np.float32), (nq,)))
    seeds = np.ascontiguousarray(random_seeds, dtype=np.uint32)
    if len(seeds) != nq:
        raise ValueError("need one seed per query")
    success = np.zeros(nq, np.int32)
    q4, t3 = np.zeros((nq, 4)), np.zeros((nq, 3))
    ninl = np.zeros(nq, np.int32)
    inl = np.zeros(max(total, 1), np.int32)
    res = np.zeros(nq, _capi.LOCALIZE_RESULT)
    rc = _load().rlbh_localize_batch(device, nq, _capi.ptr(off), _capi.ptr(p2), _capi.ptr(p3), _capi.ptr(fx),
                                     _capi.ptr(fy), C.c_float(inlier_threshold), int(number_lo_steps),
                                     int(min_num_iterations), int(max_num_iterations), _capi.ptr(seeds),
                                     _capi.ptr(success), _capi.ptr(q4), _capi.ptr(t3), _capi.ptr(ninl),
                                     _capi.ptr(inl), _capi.ptr(res))
    _check(rc)
    out = []
    for q in range(nq):
        if not success[q]:
            out.append({"success": False})
            continue
        a = int(off[q])
        out.append({"success": True, "qvec": q4[q].copy(), "tvec": t3[q].copy(), "num_inliers": int(ninl[q]),
                    "inliers": inl[a:a + ninl[q]].tolist(),
                    "_stats": {k: res[q][k].item() for k in res.dtype.names}})
    return out
