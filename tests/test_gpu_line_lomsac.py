"""Device-resident line LO-MSAC (csrc/line_lomsac_kernel.cu; SURVEY 8(f) rank 1: the covariance
line fit and with it the whole LO chain on the device) against the oracle's sequential LO-MSAC
(bit-identical to the reference's own RansacLib/ransac.h, tests/test_oracle_golden.py) on the same
seeds: iteration count, LO count, inlier index set and inlier ratio exact, score / model to 1e-9
(north_star). The kernel runs the reference's formulation without FMA contraction, so in practice
models come out bit-identical."""
import ctypes as C
import os

import numpy as np
import pytest

import oracle_lib as ol
from ransaclib_b200 import synth

pytestmark = pytest.mark.gpu

ORACLE_KEYS = ("min_iters", "max_iters", "success_prob", "sq_thr", "lo_steps", "thr_mult", "lsq_iters",
               "min_sample_mult", "non_min_sample_mult", "lo_start", "final_lsq")


@pytest.fixture(scope="module")
def ctx():
    import ransaclib_b200 as rlb

    c = rlb.Context(0)
    yield c
    c.close()


def oracle_line(xy, seed, init=None, **kw):
    O = ol.oracle()
    o = ol.Options.default(seed=int(seed), **{k: v for k, v in kw.items() if k in ORACLE_KEYS})
    st = ol.Stats()
    model = np.zeros(3) if init is None else np.array(init, dtype=np.float64)
    inl = np.zeros(max(len(xy), 1), np.int32)
    O.lomsac_line(C.byref(o), len(xy), ol.c64(xy), model, C.byref(st), inl)
    return model, st, inl[:st.best_num_inliers].copy()


def make_problems(specs, thr=0.005):
    xs = [synth.line_instance(s, n_in, n_out, thr) if n_in + n_out > 0 else np.zeros((0, 2)) for s, n_in, n_out in specs]
    off = np.concatenate([[0], np.cumsum([len(x) for x in xs])]).astype(np.int64)
    return np.ascontiguousarray(np.concatenate(xs)), off


def check_against_oracle(xy, off, seeds, lines, res, inl, init=None, **kw):
    for p in range(len(off) - 1):
        a, b = int(off[p]), int(off[p + 1])
        tag = f"problem {p} (n={b - a}, seed={seeds[p]})"
        if b - a < 2:  # ransac.h:135-139: nothing to estimate, model untouched
            assert res["success"][p] == 0, tag
            assert lines[p].tolist() == ([0.0] * 3 if init is None else list(init[p])), tag
            continue
        om, ost, oinl = oracle_line(xy[a:b], seeds[p], None if init is None else init[p], **kw)
        assert res["success"][p] == 1, tag
        assert res["num_iterations"][p] == ost.num_iterations, tag
        assert res["num_lo"][p] == ost.number_lo_iterations, tag
        assert res["num_inliers"][p] == ost.best_num_inliers, tag
        assert inl[a:a + res["num_inliers"][p]].tolist() == oinl.tolist(), tag
        np.testing.assert_allclose(res["score"][p], ost.best_model_score, rtol=1e-9, err_msg=tag)
        np.testing.assert_allclose(lines[p], om, rtol=1e-9, atol=1e-12, err_msg=tag)
        assert res["inlier_ratio"][p] == ost.inlier_ratio, tag


@pytest.mark.parametrize("outlier_ratio,seed", [(0.1, 42), (0.5, 7), (0.9, 3), (0.95, 5), (0.99, 11), (0.999, 13)])
def test_c1_line_estimation_in_one_launch(ctx, outlier_ratio, seed):
    """BASELINE configs[0] (examples/line_estimation.cc: N = 10,000, thr^2 = 1e-4, 100 .. 100k iterations)
    over a resident line scorer: one launch, same run as the reference."""
    n = 10000
    n_out = int(n * outlier_ratio)
    xy = synth.line_instance(1000 + seed, n - n_out, n_out)
    kw = dict(sq_thr=1e-4, min_iters=100, max_iters=100000)
    om, ost, oinl = oracle_line(xy, seed, **kw)
    sc = ctx.line_scorer(xy)
    before = ctx.launch_count
    line, r, inl = sc.line_lomsac(ctx.lomsac_options(**kw), seed)
    assert ctx.launch_count - before == 1
    assert r["success"] == 1 and r["num_iterations"] == ost.num_iterations
    assert r["num_lo"] == ost.number_lo_iterations and r["num_inliers"] == ost.best_num_inliers
    assert inl.tolist() == oinl.tolist()
    np.testing.assert_allclose(r["score"], ost.best_model_score, rtol=1e-9)
    np.testing.assert_allclose(line, om, rtol=1e-9, atol=1e-12)
    assert r["inlier_ratio"] == ost.inlier_ratio
    sc.close()


def test_ragged_and_degenerate_problems(ctx):
    """Empty and 1-point problems, the sampler's shuffle path (n = 2, 3), fewer points than the
    non-minimal sample (n < 6: LocalOptimization returns at once), sizes around the 512-thread CTA and
    the 16,384-point compaction round, outlier ratios 0 .. 0.95, no consensus at all."""
    specs = [(11, 0, 0), (12, 1, 0), (13, 2, 0), (14, 3, 0), (15, 4, 0), (16, 5, 0), (17, 6, 0), (18, 7, 3),
             (19, 31, 2), (20, 300, 212), (21, 400, 113), (22, 100, 1900), (23, 9000, 7384), (24, 9000, 7385),
             (25, 20000, 13001), (26, 0, 50), (27, 64, 0), (28, 2, 30), (29, 9000, 6872), (30, 9000, 6873),
             (31, 500, 12), (32, 500, 13)]
    xy, off = make_problems(specs)
    seeds = np.array([5 + 3 * i for i in range(len(specs))], np.uint32)
    kw = dict(sq_thr=1e-4, max_iters=3000)
    lines, res, inl = ctx.line_lomsac_batch(off, xy, seeds, ctx.lomsac_options(**kw))
    check_against_oracle(xy, off, seeds, lines, res, inl, **kw)


@pytest.mark.parametrize("kw", [dict(lo_steps=0), dict(final_lsq=1), dict(lo_start=0), dict(lo_start=100000),
                                dict(min_iters=700, max_iters=700), dict(lsq_iters=1), dict(lsq_iters=0),
                                dict(min_sample_mult=1), dict(min_sample_mult=16, non_min_sample_mult=16),
                                dict(thr_mult=3.0, lo_steps=3, min_iters=10), dict(success_prob=0.5, min_iters=1)])
def test_option_variants(ctx, kw):
    """Every LORansacOptions field: no LO steps, final least squares, LO from the
    first iteration / only after the loop (ransac.h:245-255), forced iteration counts, the
    (lsq_iters - 1) division of ransac.h:395-397 at 1 and 0, sample multipliers at both ends."""
    specs = [(300 + q, 600 - 80 * q, 200 + 150 * q) for q in range(6)]
    xy, off = make_problems(specs)
    seeds = np.arange(1, 1 + len(specs), dtype=np.uint32)
    kw = dict(sq_thr=1e-4, **kw)
    lines, res, inl = ctx.line_lomsac_batch(off, xy, seeds, ctx.lomsac_options(**kw))
    check_against_oracle(xy, off, seeds, lines, res, inl, **kw)


def test_callers_model_is_read_and_kept(ctx):
    """EstimateModel leaves *best_model untouched when no hypothesis is found, and its final
    least-squares step evaluates that very model (ransac.h:257-272): with every point identical the
    2-point solver never returns a line, so the result is decided by the caller's model."""
    n = 50
    xy = np.tile(np.array([[0.25, -1.5]]), (n, 1))
    off = np.array([0, n, 2 * n], np.int64)
    xy2 = np.concatenate([xy, xy])
    init = np.array([[0.6, 0.8, 1.05], [0.0, 1.0, 7.0]])  # the first passes through the point, the second does not
    seeds = np.array([3, 4], np.uint32)
    for final_lsq in (0, 1):
        kw = dict(sq_thr=1e-4, max_iters=300, final_lsq=final_lsq)
        lines, res, inl = ctx.line_lomsac_batch(off, xy2, seeds, ctx.lomsac_options(**kw), lines=init)
        check_against_oracle(xy2, off, seeds, lines, res, inl, init=init, **kw)
        assert lines.tolist() == init.tolist()
        assert res["num_inliers"].tolist() == ([n, 0] if final_lsq else [0, 0])


def test_sequential_shuffle_path_is_identical(ctx):
    """debug bit 0 sends every shuffle step through the sequential Lemire path that normally only
    handles a rejected draw (probability ~ n / 2^32 per step): nothing may change."""
    specs = [(700 + q, 300, 150) for q in range(4)]
    xy, off = make_problems(specs)
    seeds = np.arange(100, 100 + len(specs), dtype=np.uint32)
    a = ctx.line_lomsac_batch(off, xy, seeds, ctx.lomsac_options(sq_thr=1e-4))
    b = ctx.line_lomsac_batch(off, xy, seeds, ctx.lomsac_options(sq_thr=1e-4, debug_flags=1))
    assert a[0].tobytes() == b[0].tobytes() and a[1].tobytes() == b[1].tobytes()
    for p in range(len(specs)):
        lo, cnt = int(off[p]), int(a[1]["num_inliers"][p])
        assert cnt > 0 and a[2][lo:lo + cnt].tolist() == b[2][lo:lo + cnt].tolist()


def test_many_problems_deterministic_and_independent(ctx):
    """More problems than resident CTAs: every result is independent of its neighbours and of
    scheduling (same bytes as the run of the 8 distinct problems alone)."""
    base = [(900 + q, 700, 500) for q in range(8)]
    xy8, off8 = make_problems(base)
    seeds8 = np.arange(20, 28, dtype=np.uint32)
    opt = ctx.lomsac_options(sq_thr=1e-4)
    l8, r8, i8 = ctx.line_lomsac_batch(off8, xy8, seeds8, opt)
    reps = 80  # 640 problems > 2 x 148 CTAs
    xy = np.concatenate([xy8] * reps)
    off = np.concatenate([[0], np.cumsum(np.tile(np.diff(off8), reps))]).astype(np.int64)
    seeds = np.tile(seeds8, reps)
    lines, res, inl = ctx.line_lomsac_batch(off, xy, seeds, opt)
    assert lines.tobytes() == np.tile(l8, (reps, 1)).tobytes()
    assert res.tobytes() == np.tile(r8, reps).tobytes()
    for p in (0, 137, 639):
        cnt = int(res["num_inliers"][p])
        a = int(off[p])
        b = int(off8[p % 8])
        assert inl[a:a + cnt].tolist() == i8[b:b + cnt].tolist()


def test_drop_in_layer_uses_it_and_agrees_with_the_host_driven_loop():
    """LocallyOptimizedMSAC<., ., LineEstimator>::EstimateModel (include/RansacLib/ransac.h) hands the
    whole run to the kernel; RLB_HOST_DRIVER=1 keeps the one-launch-per-pass loop. Same run either way."""
    from ransaclib_b200 import host as H

    xy = synth.line_instance(1007, 5000, 5000)
    kw = dict(sq_thr=1e-4, min_iters=100, max_iters=100000)
    om, ost, oinl = oracle_line(xy, 7, **kw)
    opts = H.LORansacOptions(squared_inlier_threshold_=1e-4, random_seed_=7, min_num_iterations_=100,
                             max_num_iterations_=100000)
    runs = {}
    for mode in ("0", "1"):
        os.environ["RLB_HOST_DRIVER"] = mode
        try:
            runs[mode] = H.estimate_line(opts, xy)
        finally:
            os.environ.pop("RLB_HOST_DRIVER", None)
    for mode, (m, st) in runs.items():
        assert st["num_iterations"] == ost.num_iterations and st["number_lo_iterations"] == ost.number_lo_iterations
        assert st["inlier_indices"].tolist() == oinl.tolist(), mode
        np.testing.assert_allclose(st["best_model_score"], ost.best_model_score, rtol=1e-9)
        np.testing.assert_allclose(m, om, rtol=1e-9, atol=1e-12)
    assert runs["0"][1]["batches"] == 0 and runs["1"][1]["batches"] > 0  # which driver ran


def test_long_runs_are_handed_back_when_allowed(ctx):
    """debug bit 2 (set by the drop-in layer): a run that still needs thousands of hypotheses after 384
    iterations returns success = 2 with the model untouched, and LocallyOptimizedMSAC restarts it in the
    host-driven loop (whole-GPU sweeps): same result as the oracle, and as the kernel left alone."""
    from ransaclib_b200 import host as H

    xy = synth.line_instance(1011, 100, 9900)  # 1 % inliers: ~10,000 iterations
    kw = dict(sq_thr=1e-4, min_iters=100, max_iters=100000)
    om, ost, oinl = oracle_line(xy, 11, **kw)
    assert ost.num_iterations > 5000
    sc = ctx.line_scorer(xy)
    init = np.array([1.0, 2.0, 3.0])
    line, r, _ = sc.line_lomsac(ctx.lomsac_options(handback_after=384, handback_min_remaining=3000, **kw), 11, line=init)
    assert r["success"] == 2 and line.tolist() == init.tolist() and 384 <= r["num_iterations"] < 1024
    # ... with its state (rlb_lomsac_resume_record): magic, the iteration, the LO engine's position
    assert r["num_inliers"] == 644 and int(_[0]) == 0x524c4231 and int(_[1]) == r["num_iterations"] and 0 <= _[3] <= 624
    line, r, inl = sc.line_lomsac(ctx.lomsac_options(**kw), 11)  # not allowed to hand back: runs to the end
    assert r["success"] == 1 and r["num_iterations"] == ost.num_iterations and inl.tolist() == oinl.tolist()
    sc.close()
    opts = H.LORansacOptions(squared_inlier_threshold_=1e-4, random_seed_=11, min_num_iterations_=100,
                             max_num_iterations_=100000)
    m, st = H.estimate_line(opts, xy)
    assert st["batches"] > 0  # the host-driven loop finished it
    assert st["handbacks"] == 1 and st["handbacks_resumed"] == 1  # from where the device left off, not from scratch
    assert st["num_iterations"] == ost.num_iterations and st["inlier_indices"].tolist() == oinl.tolist()
    assert st["number_lo_iterations"] == ost.number_lo_iterations
    np.testing.assert_allclose(m, om, rtol=1e-9, atol=1e-12)
    np.testing.assert_allclose(st["best_model_score"], ost.best_model_score, rtol=1e-9)
    # hand-backs at other points of the run (before / after the first local optimisation, LO start inside the
    # device part) resume to the same run
    for seed, lo_start in ((3, 50), (5, 500), (8, 384), (13, 20)):
        om, ost, oinl = oracle_line(xy, seed, lo_start=lo_start, **kw)
        o2 = H.LORansacOptions(squared_inlier_threshold_=1e-4, random_seed_=seed, min_num_iterations_=100,
                               max_num_iterations_=100000, lo_starting_iterations_=lo_start)
        m, st = H.estimate_line(o2, xy)
        assert st["handbacks_resumed"] == 1, (seed, lo_start)
        assert st["num_iterations"] == ost.num_iterations and st["number_lo_iterations"] == ost.number_lo_iterations
        assert st["inlier_indices"].tolist() == oinl.tolist()
        np.testing.assert_allclose(m, om, rtol=1e-9, atol=1e-12)


def test_large_problems_stay_with_the_host_driven_loop():
    """Above LineEstimator::kDeviceLomsacMaxPoints (24,576) the drop-in layer keeps its own loop (whole-GPU
    passes beat one SM there); the run is the oracle's either way."""
    from ransaclib_b200 import host as H

    xy = synth.line_instance(1021, 40000, 30000)
    kw = dict(sq_thr=1e-4, min_iters=100, max_iters=100000)
    om, ost, oinl = oracle_line(xy, 5, **kw)
    opts = H.LORansacOptions(squared_inlier_threshold_=1e-4, random_seed_=5, min_num_iterations_=100,
                             max_num_iterations_=100000)
    m, st = H.estimate_line(opts, xy)
    assert st["batches"] > 0
    assert st["num_iterations"] == ost.num_iterations and st["inlier_indices"].tolist() == oinl.tolist()
    np.testing.assert_allclose(m, om, rtol=1e-9, atol=1e-12)


def test_without_inlier_output(ctx):
    """inliers = NULL: the best model's list lives in scratch; models, counts and scores do not change."""
    xy, off = make_problems([(40 + q, 500, 300) for q in range(5)])
    seeds = np.arange(9, 14, dtype=np.uint32)
    opt = ctx.lomsac_options(sq_thr=1e-4, final_lsq=1)
    a = ctx.line_lomsac_batch(off, xy, seeds, opt)
    b = ctx.line_lomsac_batch(off, xy, seeds, opt, want_inliers=False)
    assert b[2] is None and a[0].tobytes() == b[0].tobytes() and a[1].tobytes() == b[1].tobytes()


def test_bad_options_fail_loudly(ctx):
    from ransaclib_b200 import _capi

    xy, off = make_problems([(1, 50, 50)])
    seeds = np.array([1], np.uint32)
    for kw in (dict(min_sample_mult=17), dict(non_min_sample_mult=17), dict(min_sample_mult=0), dict(lo_steps=-1)):
        with pytest.raises(_capi.RlbError):
            ctx.line_lomsac_batch(off, xy, seeds, ctx.lomsac_options(**kw))
    pose = synth.pose_instance(1, 20, 20)
    sc = ctx.pose_scorer(pose["points2D"], pose["points3D"], pose["focal"], pose["focal"])
    with pytest.raises(_capi.RlbError):
        sc.line_lomsac(ctx.lomsac_options(), 1)
    sc.close()
