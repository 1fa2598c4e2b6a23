"""ransaclib_b200 — B200-native MSAC hypothesis scoring behind the RansacLib API.

Python mirror of the drop-in surface. Everything numerical happens in hand-written
CUDA kernels (``csrc/``) reached through the C ABI of ``include/ransaclib_b200.h``;
this package only marshals numpy buffers. No CPU fallback exists.
"""
from __future__ import annotations

import atexit
import ctypes as C
import weakref

import numpy as np

from . import _capi
from ._capi import (KIND_HYBRID_LINE, KIND_LINE, KIND_POSE, RLB_F32, RLB_F32_SORTED, RLB_F32_UNSORTED, RLB_F64,
                    RLB_F64_FILTERED, RLB_F64_PLAIN, RlbError)

__all__ = ["Context", "Scorer", "Comm", "CommScorer", "RLB_F64", "RLB_F32", "RLB_F64_PLAIN", "RLB_F64_FILTERED", "RLB_F32_SORTED", "RLB_F32_UNSORTED", "RlbError", "KIND_LINE", "KIND_HYBRID_LINE", "KIND_POSE"]

DBL_MAX = float(np.finfo(np.float64).max)


class Context:
    """One device + one stream + scratch memory (``rlb_ctx``)."""

    def __init__(self, device: int = 0, _borrowed=None):
        self._lib = _capi.load()
        self._owned = _borrowed is None
        if _borrowed is None:
            h = C.c_void_p()
            _capi.check(self._lib.rlb_ctx_create(device, C.byref(h)))
        else:
            h = C.c_void_p(_borrowed)  # a context owned by an rlb_comm
        self._h = h
        self.device = device
        self._scorers = weakref.WeakSet()  # scorers must be released before their context

    def close(self):
        if getattr(self, "_h", None):
            for sc in list(self._scorers):
                sc.close()
            if self._owned:
                self._lib.rlb_ctx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_stream(self, cuda_stream: int | None):
        """Run on the given cudaStream_t handle (0 = legacy default stream); None = own stream."""
        if cuda_stream is None:
            _capi.check(self._lib.rlb_ctx_use_own_stream(self._h))
        else:
            _capi.check(self._lib.rlb_ctx_set_stream(self._h, C.c_void_p(cuda_stream)))

    def synchronize(self):
        _capi.check(self._lib.rlb_ctx_synchronize(self._h))

    @property
    def launch_count(self) -> int:
        return int(self._lib.rlb_ctx_launch_count(self._h))

    @property
    def sm_count(self) -> int:
        return int(self._lib.rlb_ctx_sm_count(self._h))

    def set_timing(self, enabled: bool):
        _capi.check(self._lib.rlb_ctx_set_timing(self._h, int(enabled)))

    def last_sweep_plan(self):
        """-> (variant, splits) of the last K x N sweep (9 = tensor-core sweep)."""
        v, sp = C.c_int32(), C.c_int32()
        _capi.check(self._lib.rlb_ctx_last_sweep_plan(self._h, C.byref(v), C.byref(sp)))
        return v.value, sp.value

    def kernel_time_ms(self, which: int) -> float:
        """Device time of the last sweep (0) / inlier (1) kernel; needs set_timing(True)."""
        out = C.c_double()
        _capi.check(self._lib.rlb_ctx_kernel_time_ms(self._h, which, C.byref(out)))
        return out.value

    def select_winner_dev(self, d_records: int, world: int, record_doubles: int, d_winner: int):
        _capi.check(self._lib.rlb_select_winner_dev(self._h, C.c_void_p(d_records), world, record_doubles,
                                                    C.c_void_p(d_winner)))

    def measure_fma_peak(self, dtype: int) -> float:
        """TFLOP/s of the DFMA (0), FFMA (1) or packed FFMA2 (2) pipe, register-resident."""
        out = C.c_double()
        _capi.check(self._lib.rlb_measure_fma_peak(self._h, dtype, C.byref(out)))
        return out.value

    # -- device-resident batched localization ---------------------------------
    def localize_options(self, inlier_threshold: float, number_lo_steps: int, min_num_iterations: int,
                         max_num_iterations: int) -> "_capi.LocalizeOptions":
        """ransaclib_localization's fixed options (ransaclib_localization.cc:100-107)."""
        o = _capi.LocalizeOptions()
        self._lib.rlb_localize_default_options(C.byref(o), float(np.float32(inlier_threshold)), int(number_lo_steps),
                                               int(min_num_iterations), int(max_num_iterations))
        return o

    def localize_batch(self, offsets, points2D, points3D, fx, fy, seeds, options, want_inliers=True):
        """Q independent P3P LO-MSAC runs in one launch (host buffers; upload pipelined with the
        kernel). -> (poses [Q,12] column-major [R|c], results record array, inliers or None)."""
        off = np.ascontiguousarray(offsets, dtype=np.int64)
        nq = len(off) - 1
        p2 = _capi.f64a(points2D, (-1, 2))
        p3 = _capi.f64a(points3D, (-1, 3))
        if len(p2) != len(p3) or len(p2) != off[-1]:
            raise ValueError("points2D / points3D / offsets disagree")
        fx = np.ascontiguousarray(np.broadcast_to(np.asarray(fx, np.float64), (nq,)))
        fy = np.ascontiguousarray(np.broadcast_to(np.asarray(fy, np.float64), (nq,)))
        seeds = np.ascontiguousarray(seeds, dtype=np.uint32)
        if len(seeds) != nq:
            raise ValueError("need one seed per query")
        poses = np.zeros((nq, 12))
        res = np.zeros(nq, _capi.LOCALIZE_RESULT)
        inl = np.zeros(max(int(off[-1]), 1), np.int32) if want_inliers else None
        _capi.check(self._lib.rlb_localize_batch(self._h, nq, _capi.ptr(off), _capi.ptr(p2), _capi.ptr(p3),
                                                 _capi.ptr(fx), _capi.ptr(fy), _capi.ptr(seeds), C.byref(options),
                                                 _capi.ptr(poses), _capi.ptr(res), _capi.ptr(inl)))
        return poses, res, inl

    def localize_batch_dev(self, nq: int, max_matches: int, d_offsets: int, d_p2d: int, d_p3d: int, d_fx: int,
                           d_fy: int, d_seeds: int, options, d_poses: int, d_results: int, d_inliers: int = 0):
        """Device-pointer variant: enqueues the persistent kernel on the context stream."""
        _capi.check(self._lib.rlb_localize_batch_dev(
            self._h, nq, max_matches, C.c_void_p(d_offsets), C.c_void_p(d_p2d), C.c_void_p(d_p3d),
            C.c_void_p(d_fx), C.c_void_p(d_fy), C.c_void_p(d_seeds), C.byref(options), C.c_void_p(d_poses),
            C.c_void_p(d_results), C.c_void_p(d_inliers or 0)))

    # -- device-resident line LO-MSAC ------------------------------------------
    def lomsac_options(self, **kw) -> "_capi.LocalizeOptions":
        """LORansacOptions defaults (ransac.h:47-92); keyword names are the struct's fields
        (min_iters, max_iters, success_prob, sq_thr, lo_steps, thr_mult, lsq_iters, min_sample_mult,
        non_min_sample_mult, lo_start, final_lsq, debug_flags, handback_after, handback_min_remaining)."""
        o = _capi.LocalizeOptions()
        self._lib.rlb_lomsac_default_options(C.byref(o))
        for k, v in kw.items():
            if k not in dict(o._fields_):
                raise AttributeError(k)
            setattr(o, k, v)
        return o

    def line_lomsac_batch(self, offsets, xy, seeds, options, lines=None, want_inliers=True):
        """P independent LO-MSAC runs of the 2D line solver in one launch, one problem per CTA (host
        buffers). -> (lines [P,3], results record array, inliers or None). `lines` = the callers'
        initial models (default zeros), returned untouched for problems where nothing was found."""
        off = np.ascontiguousarray(offsets, dtype=np.int64)
        npb = len(off) - 1
        xy = _capi.f64a(xy, (-1, 2))
        if len(xy) != off[-1]:
            raise ValueError("xy / offsets disagree")
        seeds = np.ascontiguousarray(seeds, dtype=np.uint32)
        if len(seeds) != npb:
            raise ValueError("need one seed per problem")
        out = np.zeros((npb, 3)) if lines is None else np.array(_capi.f64a(lines, (npb, 3)), copy=True)
        res = np.zeros(npb, _capi.LOCALIZE_RESULT)
        inl = np.zeros(max(int(off[-1]), 1), np.int32) if want_inliers else None
        _capi.check(self._lib.rlb_line_lomsac_batch(self._h, npb, _capi.ptr(off), _capi.ptr(xy), _capi.ptr(seeds),
                                                    C.byref(options), _capi.ptr(out), _capi.ptr(res), _capi.ptr(inl)))
        return out, res, inl

    def line_lomsac_batch_dev(self, num_problems: int, max_points: int, d_offsets: int, d_xy: int, d_seeds: int,
                              options, d_lines: int, d_results: int, d_inliers: int = 0):
        """Device-pointer variant: enqueues the kernel on the context stream."""
        _capi.check(self._lib.rlb_line_lomsac_batch_dev(
            self._h, num_problems, max_points, C.c_void_p(d_offsets), C.c_void_p(d_xy), C.c_void_p(d_seeds),
            C.byref(options), C.c_void_p(d_lines), C.c_void_p(d_results), C.c_void_p(d_inliers or 0)))

    # -- data sets -----------------------------------------------------------
    def line_scorer(self, xy) -> "Scorer":
        xy = _capi.f64a(xy, (-1, 2))
        h = C.c_void_p()
        _capi.check(self._lib.rlb_scorer_create_line(self._h, len(xy), _capi.ptr(xy), C.byref(h)))
        return Scorer(self, h)

    def hybrid_line_scorer(self, xy0, xynn1) -> "Scorer":
        xy0 = _capi.f64a(xy0, (-1, 2))
        xynn1 = _capi.f64a(xynn1, (-1, 4))
        h = C.c_void_p()
        _capi.check(self._lib.rlb_scorer_create_hybrid_line(self._h, len(xy0), _capi.ptr(xy0), len(xynn1),
                                                            _capi.ptr(xynn1), C.byref(h)))
        return Scorer(self, h)

    def custom_scorer(self, data, model_doubles: int, eval_source: str) -> "Scorer":
        """Any other solver of the reference's concept: ``data`` is n x dims (one point per row) and
        ``eval_source`` is CUDA source defining
        ``__device__ double rlb_eval(const double* model, const double* point)`` = the squared residual
        (the body of EvaluateModelOnPoint). Compiled with NVRTC at creation."""
        data = np.ascontiguousarray(data, dtype=np.float64)
        if data.ndim != 2:
            raise ValueError("data must be n x dims")
        h = C.c_void_p()
        _capi.check(self._lib.rlb_scorer_create_custom(self._h, data.shape[0], data.shape[1], _capi.ptr(data),
                                                       int(model_doubles), eval_source.encode(), C.byref(h)))
        return Scorer(self, h)

    def custom_hybrid_scorer(self, data0, data1, model_doubles: int, eval_source: str) -> "Scorer":
        """The HybridSolver form of a user-defined functor (``rlb_scorer_create_custom_hybrid``): two data types,
        ``data_t`` is n_t x dims_t, ``eval_source`` defines
        ``__device__ double rlb_eval_t(const double* model, int type, const double* point)``; thresholds and
        weights are per type as for the built-in hybrid line scorer."""
        d0 = np.ascontiguousarray(data0, dtype=np.float64)
        d1 = np.ascontiguousarray(data1, dtype=np.float64)
        if d0.ndim != 2 or d1.ndim != 2:
            raise ValueError("data0 / data1 must be n x dims")
        h = C.c_void_p()
        _capi.check(self._lib.rlb_scorer_create_custom_hybrid(
            self._h, d0.shape[0], d0.shape[1], _capi.ptr(d0), d1.shape[0], d1.shape[1], _capi.ptr(d1),
            int(model_doubles), eval_source.encode(), C.byref(h)))
        return Scorer(self, h)

    def pose_scorer(self, points2D, points3D, fx: float, fy: float) -> "Scorer":
        p2 = _capi.f64a(points2D, (-1, 2))
        p3 = _capi.f64a(points3D, (-1, 3))
        if len(p2) != len(p3):
            raise ValueError("points2D and points3D must have the same length")
        h = C.c_void_p()
        _capi.check(self._lib.rlb_scorer_create_pose(self._h, len(p2), _capi.ptr(p2), _capi.ptr(p3), float(fx),
                                                     float(fy), C.byref(h)))
        return Scorer(self, h)


# Communicators still alive when the interpreter exits are destroyed (scorers first) BEFORE the CUDA runtime and
# NCCL tear themselves down: a __del__ that reaches ncclCommDestroy during interpreter shutdown can crash.
_LIVE_COMMS = weakref.WeakSet()


@atexit.register
def _close_live_comms():
    for comm in list(_LIVE_COMMS):
        try:
            comm.close()
        except Exception:
            pass


class Comm:
    """Hypothesis-sharded scoring over several GPUs with the collective inside the library (``rlb_comm``):
    ``Comm.local(ndev)`` drives ``ndev`` GPUs from this process, ``Comm.from_rank(ctx, nranks, rank, uid)`` is
    the one-process-per-GPU form (``Comm.unique_id()`` on rank 0, handed to the others by the caller)."""

    def __init__(self, handle, ctx=None):
        self._lib = _capi.load()
        self._h = handle
        self._ctx = ctx  # keeps a caller-owned context alive
        self._scorers = weakref.WeakSet()
        _LIVE_COMMS.add(self)

    @staticmethod
    def unique_id() -> bytes:
        buf = C.create_string_buffer(_capi.COMM_ID_BYTES)
        _capi.check(_capi.load().rlb_comm_get_unique_id(buf))
        return buf.raw

    @classmethod
    def from_rank(cls, ctx: "Context", nranks: int, rank: int, uid: bytes | None) -> "Comm":
        h = C.c_void_p()
        buf = C.create_string_buffer(uid, _capi.COMM_ID_BYTES) if uid is not None else None
        _capi.check(_capi.load().rlb_comm_init_rank(ctx._h, int(nranks), int(rank), buf, C.byref(h)))
        return cls(h, ctx)

    @classmethod
    def local(cls, ndev: int, devices=None) -> "Comm":
        h = C.c_void_p()
        d = None if devices is None else _capi.i32a(devices)
        _capi.check(_capi.load().rlb_comm_create_local(int(ndev), _capi.ptr(d), C.byref(h)))
        return cls(h)

    @property
    def nranks(self) -> int:
        return int(self._lib.rlb_comm_nranks(self._h))

    @property
    def num_local(self) -> int:
        return int(self._lib.rlb_comm_num_local(self._h))

    def rank(self, local_index: int = 0) -> int:
        return int(self._lib.rlb_comm_rank(self._h, local_index))

    @property
    def collective_count(self) -> int:
        return int(self._lib.rlb_comm_collective_count(self._h))

    def context(self, local_index: int = 0) -> "Context":
        """The context of a local GPU (owned by the communicator or by the caller; not closed here)."""
        return Context(0, _borrowed=self._lib.rlb_comm_ctx(self._h, local_index))

    def pose_scorer(self, points2D, points3D, fx: float, fy: float, root: int = 0) -> "CommScorer":
        """Collective. The arrays are read on the process that owns rank `root` (one PCIe upload) and reach
        the other GPUs by ncclBroadcast; other processes may pass None with the number of points.
        root = -1 (RLB_COMM_ROOT_ALL): every process holds the arrays; each GPU uploads 1/G of them over its
        own PCIe link and the slices are exchanged over NVLink."""
        if isinstance(points2D, int):
            n, p2, p3 = points2D, None, None
        else:
            p2 = _capi.f64a(points2D, (-1, 2))
            p3 = _capi.f64a(points3D, (-1, 3))
            n = len(p2)
        h = C.c_void_p()
        _capi.check(self._lib.rlb_comm_scorer_create_pose(self._h, n, _capi.ptr(p2), _capi.ptr(p3), float(fx),
                                                          float(fy), int(root), C.byref(h)))
        cs = CommScorer(self, h)
        self._scorers.add(cs)
        return cs

    def close(self):
        if getattr(self, "_h", None):
            for cs in list(self._scorers):
                cs.close()
            self._lib.rlb_comm_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class CommScorer:
    """One pose data set replicated on every GPU of a communicator (``rlb_comm_scorer``)."""

    def __init__(self, comm: Comm, handle):
        self.comm = comm
        self._lib = comm._lib
        self._h = handle
        self.model_size = 12

    def close(self):
        if getattr(self, "_h", None):
            self._lib.rlb_comm_scorer_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def hypothesize_and_score(self, samples, sq_thr: float, precision=RLB_F64, want_scores=False):
        """One sharded round over ALL the samples (k_total x 4, the same array on every process):
        -> dict(best_id (global), best_score, best_model[, scores, counts: this process's slices filled])."""
        s = _capi.i32a(samples, (-1, 4))
        k = len(s)
        thr = np.array([sq_thr], np.float64)
        scores = np.full(k, np.nan) if want_scores else None
        counts = np.full(k, -1, np.int32) if want_scores else None
        bid, bsc, bm = C.c_int64(), C.c_double(), np.zeros(12)
        _capi.check(self._lib.rlb_comm_hypothesize_and_score(
            self._h, precision, 0, k, _capi.ptr(s), _capi.ptr(thr), None, _capi.ptr(scores), _capi.ptr(counts),
            C.addressof(bid), C.addressof(bsc), _capi.ptr(bm)))
        return {"best_id": int(bid.value), "best_score": bsc.value, "best_model": bm if bid.value >= 0 else None,
                "scores": scores, "counts": counts}

    def round_dev(self, d_samples: int, k_local: int, sq_thr: float, id_offset: int, d_models: int, d_valid: int,
                  d_scores: int, d_counts: int, d_winner: int, precision=RLB_F64):
        """Device-pointer round of the one-process-per-GPU communicator (asynchronous)."""
        thr = np.array([sq_thr], np.float64)
        _capi.check(self._lib.rlb_comm_round_dev(
            self._h, precision, 0, k_local, C.c_void_p(d_samples), _capi.ptr(thr), None, int(id_offset),
            C.c_void_p(d_models), C.c_void_p(d_valid), C.c_void_p(d_scores or 0), C.c_void_p(d_counts or 0),
            C.c_void_p(d_winner)))

    def local_scorer_handle(self, local_index: int = 0) -> int:
        return self._lib.rlb_comm_scorer_local(self._h, local_index)

    def get_inliers(self, model, sq_thr_inl: float, sq_thr_score: float | None = None):
        """-> (ascending inlier indices, score or None); answered by the first local GPU."""
        model = _capi.f64a(model, (12,))
        ti = np.array([sq_thr_inl], np.float64)
        ts = None if sq_thr_score is None else np.array([sq_thr_score], np.float64)
        n = int(self._lib.rlb_scorer_num_data(C.c_void_p(self.local_scorer_handle(0)), 0))
        idx = getattr(self.comm, "_idx_buf", None)  # reused across calls (a fresh 4 MB array costs page faults)
        if idx is None or len(idx) < max(n, 1):
            idx = self.comm._idx_buf = np.empty(max(n, 1), np.int32)
        counts = np.zeros(2, np.int32)
        score = C.c_double()
        _capi.check(self._lib.rlb_comm_get_inliers(self._h, _capi.ptr(model), _capi.ptr(ti), _capi.ptr(ts), None,
                                                   _capi.ptr(idx), _capi.ptr(counts), C.addressof(score)))
        return idx[:counts[0]].copy(), (score.value if ts is not None else None)


def custom_compile_check(eval_source: str, dims: int, model_doubles: int) -> int:
    """Compiles a user functor the way ``Context.custom_scorer`` would (NVRTC, sm_100a) without touching a
    device; -> size of the cubin. A compile error raises RlbError carrying the compiler log."""
    lib = _capi.load()
    n = C.c_size_t()
    _capi.check(lib.rlb_custom_compile_check(eval_source.encode(), int(dims), int(model_doubles), C.byref(n)))
    return int(n.value)


class Scorer:
    """A data set resident in HBM plus the functor that scores models on it (``rlb_scorer``)."""

    def __init__(self, ctx: Context, handle, borrowed: bool = False):
        self.ctx = ctx
        self._lib = ctx._lib
        self._h = C.c_void_p(handle) if isinstance(handle, int) else handle
        handle = self._h
        self._borrowed = borrowed  # owned by an rlb_comm_scorer: never destroyed here
        if not borrowed:
            ctx._scorers.add(self)
        self.kind = self._lib.rlb_scorer_kind(handle)
        self.num_types = self._lib.rlb_scorer_num_types(handle)
        self.num_data = [self._lib.rlb_scorer_num_data(handle, t) for t in range(self.num_types)]
        self.model_size = self._lib.rlb_scorer_model_size(handle)

    def close(self):
        if getattr(self, "_h", None):
            if not self._borrowed:
                self._lib.rlb_scorer_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _thr(self, v):
        a = np.atleast_1d(np.asarray(v, dtype=np.float64))
        if len(a) == 1 and self.num_types == 2:
            a = np.repeat(a, 2)
        if len(a) != self.num_types:
            raise ValueError("need one value per data type")
        return np.ascontiguousarray(a)

    def rays(self):
        out = np.empty((self.num_data[0], 3))
        _capi.check(self._lib.rlb_scorer_get_rays(self._h, _capi.ptr(out)))
        return out

    def score_models(self, models, sq_thr, weights=None, valid=None, precision=RLB_F64, want_counts=True):
        """-> (scores[K], counts[K] or None, best_id, best_score). want_counts=False passes a NULL counts
        array, i.e. asks for what ScoreModel (ransac.h) computes; the sweep then skips the counting."""
        models = _capi.f64a(models, (-1, self.model_size))
        k = len(models)
        thr = self._thr(sq_thr)
        w = None if weights is None else self._thr(weights)
        v = None if valid is None else np.ascontiguousarray(valid, dtype=np.uint8)
        scores = np.empty(k)
        counts = np.empty(k, np.int32) if want_counts else None
        bid = C.c_int32()
        bsc = C.c_double()
        _capi.check(self._lib.rlb_score_models(self._h, precision, k, _capi.ptr(models), _capi.ptr(v),
                                               _capi.ptr(thr), _capi.ptr(w), _capi.ptr(scores), _capi.ptr(counts),
                                               C.addressof(bid), C.addressof(bsc)))
        return scores, counts, bid.value, bsc.value

    def get_inliers(self, model, sq_thr_inl, sq_thr_score=None, weights=None, want_indices=True):
        """-> (list of per-type ascending index arrays, per-type counts, score or None)."""
        model = _capi.f64a(model, (self.model_size,))
        ti = self._thr(sq_thr_inl)
        ts = None if sq_thr_score is None else self._thr(sq_thr_score)
        w = None if weights is None else self._thr(weights)
        idx = np.empty(max(sum(self.num_data), 1), np.int32) if want_indices else None
        counts = np.zeros(2, np.int32)
        score = C.c_double()
        _capi.check(self._lib.rlb_get_inliers(self._h, _capi.ptr(model), _capi.ptr(ti), _capi.ptr(ts), _capi.ptr(w),
                                              _capi.ptr(idx), _capi.ptr(counts), C.addressof(score)))
        lists = None
        if want_indices:
            lists, p = [], 0
            for t in range(self.num_types):
                lists.append(idx[p:p + counts[t]].copy())
                p += counts[t]
        return lists, counts[:self.num_types].copy(), (score.value if ts is not None else None)

    def get_inliers_view(self, model, sq_thr_inl, sq_thr_score=None, weights=None):
        """The same pass through rlb_get_inliers_view: the lists are read from the context's own pinned copy
        (valid until the next call on the context; copied here before returning)."""
        model = _capi.f64a(model, (self.model_size,))
        ti = self._thr(sq_thr_inl)
        ts = None if sq_thr_score is None else self._thr(sq_thr_score)
        w = None if weights is None else self._thr(weights)
        view = C.c_void_p()
        counts = np.zeros(2, np.int32)
        score = C.c_double()
        _capi.check(self._lib.rlb_get_inliers_view(self._h, _capi.ptr(model), _capi.ptr(ti), _capi.ptr(ts),
                                                   _capi.ptr(w), C.byref(view), _capi.ptr(counts),
                                                   C.addressof(score)))
        total = int(counts[:self.num_types].sum())
        flat = np.empty(0, np.int32)
        if total > 0:
            flat = np.ctypeslib.as_array(C.cast(view, C.POINTER(C.c_int32)), shape=(total,)).copy()
        lists, p = [], 0
        for t in range(self.num_types):
            lists.append(flat[p:p + counts[t]])
            p += counts[t]
        return lists, counts[:self.num_types].copy(), (score.value if ts is not None else None)

    def eval_errors(self, model, type_: int = 0, precision=RLB_F64):
        model = _capi.f64a(model, (self.model_size,))
        err = np.empty(self.num_data[type_])
        _capi.check(self._lib.rlb_eval_errors(self._h, precision, _capi.ptr(model), type_, _capi.ptr(err)))
        return err

    def line_lomsac(self, options, seed: int, line=None, want_inliers=True):
        """Whole LO-MSAC run over this LINE scorer's resident data in one launch (one CTA).
        -> (line[3], result record, ascending inlier indices or None)."""
        out = np.zeros(3) if line is None else np.array(_capi.f64a(line, (3,)), copy=True)
        res = np.zeros(1, _capi.LOCALIZE_RESULT)
        n = self.num_data[0]
        inl = np.zeros(max(n, 1), np.int32) if want_inliers else None
        _capi.check(self._lib.rlb_scorer_line_lomsac(self._h, C.byref(options), int(seed), _capi.ptr(out),
                                                     _capi.ptr(res), _capi.ptr(inl)))
        r = res[0]
        return out, r, (inl[:int(r["num_inliers"])].copy() if want_inliers else None)

    def solve_line2pt(self, samples):
        s = _capi.i32a(samples, (-1, 2))
        models = np.empty((len(s), 3))
        valid = np.empty(len(s), np.uint8)
        _capi.check(self._lib.rlb_solve_line2pt_batch(self._h, len(s), _capi.ptr(s), _capi.ptr(models),
                                                      _capi.ptr(valid)))
        return models, valid

    def solve_point_normal(self, samples):
        s = _capi.i32a(samples, (-1,))
        models = np.empty((len(s), 3))
        valid = np.empty(len(s), np.uint8)
        _capi.check(self._lib.rlb_solve_point_normal_batch(self._h, len(s), _capi.ptr(s), _capi.ptr(models),
                                                           _capi.ptr(valid)))
        return models, valid

    def solve_p3p(self, samples, sq_thr: float):
        s = _capi.i32a(samples, (-1, 4))
        models = np.empty((len(s), 12))
        valid = np.empty(len(s), np.uint8)
        _capi.check(self._lib.rlb_solve_p3p_batch(self._h, len(s), _capi.ptr(s), float(sq_thr), _capi.ptr(models),
                                                  _capi.ptr(valid)))
        return models, valid

    # -- device-pointer entry points (inputs resident in HBM; asynchronous) ----
    def hypothesize_and_score_dev(self, d_samples: int, k: int, sq_thr, weights, d_models: int, d_valid: int,
                                  d_scores: int, d_counts: int, d_best: int, solver: int = 0, precision=RLB_F64):
        thr = self._thr(sq_thr)
        w = None if weights is None else self._thr(weights)
        _capi.check(self._lib.rlb_hypothesize_and_score_dev(
            self._h, precision, solver, k, C.c_void_p(d_samples), _capi.ptr(thr), _capi.ptr(w),
            C.c_void_p(d_models), C.c_void_p(d_valid), C.c_void_p(d_scores or 0), C.c_void_p(d_counts or 0),
            C.c_void_p(d_best)))

    def sharded_round_dev(self, d_samples: int, k: int, sq_thr, weights, id_offset: int, d_models: int,
                          d_valid: int, d_scores: int, d_counts: int, d_record: int, solver: int = 0,
                          precision=RLB_F64):
        """One hypothesis-sharded round: solve + sweep + argmin, local winner record -> d_record."""
        thr = self._thr(sq_thr)
        w = None if weights is None else self._thr(weights)
        _capi.check(self._lib.rlb_sharded_round_dev(
            self._h, precision, solver, k, C.c_void_p(d_samples), _capi.ptr(thr), _capi.ptr(w), int(id_offset),
            C.c_void_p(d_models), C.c_void_p(d_valid), C.c_void_p(d_scores or 0), C.c_void_p(d_counts or 0),
            C.c_void_p(d_record)))

    def score_models_dev(self, d_models: int, k: int, sq_thr, weights, d_valid: int, d_scores: int, d_counts: int,
                         d_best: int, precision=RLB_F64):
        thr = self._thr(sq_thr)
        w = None if weights is None else self._thr(weights)
        _capi.check(self._lib.rlb_score_models_dev(
            self._h, precision, k, C.c_void_p(d_models), C.c_void_p(d_valid or 0), _capi.ptr(thr), _capi.ptr(w),
            C.c_void_p(d_scores or 0), C.c_void_p(d_counts or 0), C.c_void_p(d_best)))

    def get_inliers_dev(self, d_model: int, sq_thr_inl, sq_thr_score, weights, d_idx: int, d_header: int):
        ti = self._thr(sq_thr_inl)
        ts = None if sq_thr_score is None else self._thr(sq_thr_score)
        w = None if weights is None else self._thr(weights)
        _capi.check(self._lib.rlb_get_inliers_dev(self._h, C.c_void_p(d_model), _capi.ptr(ti), _capi.ptr(ts),
                                                  _capi.ptr(w), C.c_void_p(d_idx or 0), C.c_void_p(d_header)))

    def get_inliers_noalloc_dev(self, d_model: int, sq_thr_inl, sq_thr_score, weights, d_idx: int, d_header: int):
        """get_inliers_dev through rlb_get_inliers_noalloc_dev: scratch in a fixed static device array, no
        allocation on the call path (at most _capi.NOALLOC_MAX_POINTS points; calls must not overlap on the GPU)."""
        ti = self._thr(sq_thr_inl)
        ts = None if sq_thr_score is None else self._thr(sq_thr_score)
        w = None if weights is None else self._thr(weights)
        _capi.check(self._lib.rlb_get_inliers_noalloc_dev(self._h, C.c_void_p(d_model), _capi.ptr(ti), _capi.ptr(ts),
                                                          _capi.ptr(w), C.c_void_p(d_idx or 0), C.c_void_p(d_header)))

    def hypothesize_and_score(self, samples, sq_thr, weights=None, solver: int = 0, precision=RLB_F64,
                              want_models=True, want_scores=True):
        """Samples up, hypotheses stay on the device. -> dict."""
        ss = {(KIND_POSE, 0): 4, (KIND_LINE, 0): 2, (KIND_HYBRID_LINE, 0): 2, (KIND_HYBRID_LINE, 1): 1}[
            (self.kind, solver)]
        s = _capi.i32a(samples, (-1, ss))
        k = len(s)
        thr = self._thr(sq_thr)
        w = None if weights is None else self._thr(weights)
        models = np.empty((k, self.model_size)) if want_models else None
        valid = np.empty(k, np.uint8) if want_models else None
        scores = np.empty(k) if want_scores else None
        counts = np.empty(k, np.int32) if want_scores else None
        bid = C.c_int32()
        bsc = C.c_double()
        bm = np.zeros(self.model_size)
        _capi.check(self._lib.rlb_hypothesize_and_score(
            self._h, precision, solver, k, _capi.ptr(s), _capi.ptr(thr), _capi.ptr(w), _capi.ptr(models),
            _capi.ptr(valid), _capi.ptr(scores), _capi.ptr(counts), C.addressof(bid), C.addressof(bsc),
            _capi.ptr(bm)))
        return {"models": models, "valid": valid, "scores": scores, "counts": counts, "best_id": bid.value,
                "best_score": bsc.value, "best_model": bm if bid.value >= 0 else None}
