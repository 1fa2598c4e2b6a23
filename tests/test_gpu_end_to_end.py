"""GPU end-to-end parity: the drop-in template layer (include/RansacLib/*.h, speculative
batches + sequential replay, fused LO passes) against the oracle's sequential LO-MSAC on the
same seeds. Same seed => same iteration count, LO count, inlier count, inlier index set and
model as the (pinned) oracle, which is bit-identical to the reference's own headers."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol
from ransaclib_b200 import synth

pytestmark = pytest.mark.gpu


def host():
    from ransaclib_b200 import host as h

    return h


def oracle_line(xy, **kw):
    O = ol.oracle()
    o = ol.Options.default(**kw)
    st = ol.Stats()
    model = np.zeros(3)
    inl = np.zeros(len(xy), np.int32)
    O.lomsac_line(C.byref(o), len(xy), ol.c64(xy), model, C.byref(st), inl)
    return model, st, inl[:st.best_num_inliers].copy()


def oracle_pose(p2, p3, f, **kw):
    O = ol.oracle()
    o = ol.Options.default(**kw)
    st = ol.Stats()
    model = np.zeros(12)
    inl = np.zeros(len(p2), np.int32)
    O.lomsac_pose(C.byref(o), f, f, len(p2), ol.c64(p2), ol.c64(p3), model, C.byref(st), inl)
    return model, st, inl[:st.best_num_inliers].copy()


def check_same(stats, model, ost, omodel, oinl, model_rtol=1e-9):
    assert stats["num_iterations"] == ost.num_iterations
    assert stats["number_lo_iterations"] == ost.number_lo_iterations
    assert stats["best_num_inliers"] == ost.best_num_inliers
    assert stats["inlier_indices"].tolist() == oinl.tolist()
    np.testing.assert_allclose(stats["best_model_score"], ost.best_model_score, rtol=1e-9)
    np.testing.assert_allclose(model, omodel, rtol=model_rtol, atol=1e-12)
    assert stats["inlier_ratio"] == ost.inlier_ratio


@pytest.mark.parametrize("outlier_ratio,seed", [(0.1, 42), (0.5, 7), (0.9, 3), (0.95, 5), (0.99, 11), (0.999, 13)])
def test_line_estimation_config(outlier_ratio, seed):
    """BASELINE configs[0]: examples/line_estimation.cc (N = 10,000, thr^2 = 1e-4, max 100k)."""
    H = host()
    n = 10000
    n_out = int(n * outlier_ratio)
    xy = synth.line_instance(1000 + seed, n - n_out, n_out)
    kw = dict(sq_thr=1e-4, seed=seed, min_iters=100, max_iters=100000)
    om, ost, oinl = oracle_line(xy, **kw)
    opts = H.LORansacOptions(squared_inlier_threshold_=1e-4, random_seed_=seed, min_num_iterations_=100,
                             max_num_iterations_=100000)
    m, st = H.estimate_line(opts, xy)
    check_same(st, m, ost, om, oinl)
    m32, st32 = H.estimate_line(opts, xy, precision=1)  # f32 sweep + f64 verification: same run
    check_same(st32, m32, ost, om, oinl)


def circle_instance(seed, n, outlier_ratio):
    rng = np.random.default_rng(seed)
    ang = rng.uniform(0, 2 * np.pi, n)
    pts = np.stack([1.5 + 2.0 * np.cos(ang), -0.5 + 2.0 * np.sin(ang)], 1) + rng.normal(0, 0.01, (n, 2))
    out = rng.random(n) < outlier_ratio
    pts[out] = rng.uniform(-3, 4, (int(out.sum()), 2))
    return np.ascontiguousarray(pts)


@pytest.mark.parametrize("n,outlier_ratio,seed,final_lsq", [(3000, 0.3, 42, 0), (20000, 0.6, 7, 1), (500, 0.8, 3, 0)])
def test_user_defined_solver_circle(n, outlier_ratio, seed, final_lsq):
    """a solver that is not one of the reference's examples (examples/circle_estimator.h:
    host MinimalSolver / Kasa fit, NVRTC-compiled residual) through the drop-in LocallyOptimizedMSAC, against the
    same solver driven by the sequential oracle (== the reference's unmodified ransac.h, test_oracle_golden.py)."""
    H = host()
    O = ol.oracle()
    xy = circle_instance(seed, n, outlier_ratio)
    kw = dict(sq_thr=0.03 ** 2, seed=seed, final_lsq=final_lsq, min_iters=100, max_iters=5000)
    o, ost = ol.Options.default(**kw), ol.Stats()
    om, oinl = np.zeros(3), np.zeros(n, np.int32)
    O.lomsac_circle(C.byref(o), n, ol.c64(xy), om, C.byref(ost), oinl)
    opts = H.LORansacOptions(squared_inlier_threshold_=0.03 ** 2, random_seed_=seed,
                             final_least_squares_=bool(final_lsq), min_num_iterations_=100, max_num_iterations_=5000)
    m, st = H.estimate_circle(opts, xy)
    check_same(st, m, ost, om, oinl[:ost.best_num_inliers])
    assert st["best_num_inliers"] > 0.8 * (1 - outlier_ratio) * n
    np.testing.assert_allclose(m, [1.5, -0.5, 2.0], atol=0.02)


@pytest.mark.parametrize("seed,final_lsq,lo_steps", [(42, 0, 10), (5, 1, 10), (9, 0, 0)])
def test_camera_pose_estimation_config(seed, final_lsq, lo_steps):
    """BASELINE configs[1] shape: N = 10k matches, 50 % outliers, thr = 12 px, adaptive stop."""
    H = host()
    inst = synth.pose_instance(2000 + seed, 5000, 5000)
    f = inst["focal"]
    kw = dict(sq_thr=144.0, seed=seed, final_lsq=final_lsq, lo_steps=lo_steps, min_iters=100, max_iters=10000)
    om, ost, oinl = oracle_pose(inst["points2D"], inst["points3D"], f, **kw)
    opts = H.LORansacOptions(squared_inlier_threshold_=144.0, random_seed_=seed, final_least_squares_=bool(final_lsq),
                             num_lo_steps_=lo_steps, min_num_iterations_=100, max_num_iterations_=10000)
    m, st = H.estimate_pose(opts, f, f, inst["points2D"], inst["points3D"])
    check_same(st, m, ost, om, oinl)
    assert st["best_num_inliers"] >= 4900
    # fused passes: far fewer than the reference's ~105 per LO call
    assert st["single_passes"] < 80 * max(1, st["number_lo_iterations"]) + 20
    m32, st32 = H.estimate_pose(opts, f, f, inst["points2D"], inst["points3D"], precision=1)
    check_same(st32, m32, ost, om, oinl)


def test_pose_forced_10k_hypotheses():
    """configs[1] exactly: min = max = 10,000 iterations forces all 10k hypotheses."""
    H = host()
    inst = synth.pose_instance(4242, 5000, 5000)
    f = inst["focal"]
    kw = dict(sq_thr=144.0, seed=42, min_iters=10000, max_iters=10000)
    om, ost, oinl = oracle_pose(inst["points2D"], inst["points3D"], f, **kw)
    opts = H.LORansacOptions(squared_inlier_threshold_=144.0, random_seed_=42, min_num_iterations_=10000,
                             max_num_iterations_=10000)
    m, st = H.estimate_pose(opts, f, f, inst["points2D"], inst["points3D"])
    assert st["num_iterations"] == 10000
    check_same(st, m, ost, om, oinl)


def test_hybrid_line_estimation_config():
    """BASELINE configs[2]: hybrid line, N0 = N1 = 100k (also the example's 100 + 100)."""
    H = host()
    O = ol.oracle()
    for n, seed in [(100, 42), (100000, 6)]:
        n_out = n // 2
        xy0, p1 = synth.hybrid_line_instance(300 + seed, n - n_out, n_out, n - n_out, n_out)
        o = ol.HybridOptions.default(sq_thr=[1e-4, 1e-4], weights=[2.0, 0.5], seed=seed, min_iters=100,
                                     max_iters=10000, max_iters_per_solver=1000)
        ost = ol.HybridStats()
        om = np.zeros(3)
        a, b = np.zeros(n, np.int32), np.zeros(n, np.int32)
        O.hybrid_lomsac_line(C.byref(o), n, ol.c64(xy0), n, ol.c64(p1), np.array([0.2, 0.8]), om, C.byref(ost), a, b)
        opts = H.HybridLORansacOptions(squared_inlier_thresholds_=[1e-4, 1e-4], data_type_weights_=[2.0, 0.5],
                                       random_seed_=seed, min_num_iterations_=100, max_num_iterations_=10000,
                                       max_num_iterations_per_solver_=1000)
        m, st = H.estimate_hybrid_line(opts, xy0, p1, [0.2, 0.8])
        assert st["num_iterations_total"] == ost.num_iterations_total
        assert st["num_iterations_per_solver"] == list(ost.num_iterations_per_solver)
        assert st["best_num_inliers"] == ost.best_num_inliers
        assert st["best_solver_type"] == ost.best_solver_type
        assert st["number_lo_iterations"] == ost.number_lo_iterations
        assert st["inlier_ratios"] == list(ost.inlier_ratios)
        np.testing.assert_allclose(st["best_model_score"], ost.best_model_score, rtol=1e-9)
        np.testing.assert_allclose(m, om, rtol=1e-9, atol=1e-12)
        assert st["inlier_indices"][0].tolist() == a[:ost.num_inliers_per_type[0]].tolist()
        assert st["inlier_indices"][1].tolist() == b[:ost.num_inliers_per_type[1]].tolist()


@pytest.mark.parametrize("variant", [
    dict(), dict(lo_steps=1), dict(lo_steps=3, lsq_iters=1), dict(lsq_iters=2, thr_mult=3.0), dict(thr_mult=1.0),
    dict(lsq_iters=7, thr_mult=1.1, min_sample_mult=2), dict(lo_steps=25, lo_start=10), dict(final_lsq=1),
    dict(lsq_iters=0), dict(lo_steps=0)])
def test_hybrid_lo_variants(variant):
    """The hybrid LO leaves out passes whose result is known (relaxed thresholds <= 0: empty lists; an unchanged model
    scored again) - hybrid_ransac.h:500-567 run by the oracle must see the same solver calls, RNG draws and updates
    whatever the LO options make of the shrinking thresholds."""
    H = host()
    O = ol.oracle()
    names = {"lo_steps": "num_lo_steps_", "lsq_iters": "num_lsq_iterations_", "thr_mult": "threshold_multiplier_",
             "min_sample_mult": "min_sample_multiplicator_", "lo_start": "lo_starting_iterations_",
             "final_lsq": "final_least_squares_"}
    for n, n_out, seed in [(3000, 1500, 3), (400, 300, 8), (5000, 500, 21)]:
        xy0, p1 = synth.hybrid_line_instance(700 + seed, n - n_out, n_out, n - n_out, n_out)
        o = ol.HybridOptions.default(sq_thr=[1e-4, 4e-4], weights=[2.0, 0.5], seed=seed, min_iters=100,
                                     max_iters=10000, max_iters_per_solver=1000, **variant)
        ost, om = ol.HybridStats(), np.zeros(3)
        a, b = np.zeros(n, np.int32), np.zeros(n, np.int32)
        O.hybrid_lomsac_line(C.byref(o), n, ol.c64(xy0), n, ol.c64(p1), np.array([0.2, 0.8]), om, C.byref(ost), a, b)
        kw = {names[k]: (bool(v) if k == "final_lsq" else v) for k, v in variant.items()}
        opts = H.HybridLORansacOptions(squared_inlier_thresholds_=[1e-4, 4e-4], data_type_weights_=[2.0, 0.5],
                                       random_seed_=seed, min_num_iterations_=100, max_num_iterations_=10000,
                                       max_num_iterations_per_solver_=1000, **kw)
        m, st = H.estimate_hybrid_line(opts, xy0, p1, [0.2, 0.8])
        assert st["num_iterations_total"] == ost.num_iterations_total, (variant, n)
        assert st["num_iterations_per_solver"] == list(ost.num_iterations_per_solver)
        assert st["best_num_inliers"] == ost.best_num_inliers and st["best_solver_type"] == ost.best_solver_type
        assert st["number_lo_iterations"] == ost.number_lo_iterations
        np.testing.assert_allclose(st["best_model_score"], ost.best_model_score, rtol=1e-9)
        np.testing.assert_allclose(m, om, rtol=1e-9, atol=1e-12)
        assert st["inlier_indices"][0].tolist() == a[:ost.num_inliers_per_type[0]].tolist()
        assert st["inlier_indices"][1].tolist() == b[:ost.num_inliers_per_type[1]].tolist()


def test_hybrid_line_with_biased_sampling():
    """HybridBiasedSampling (reference hybrid_sampling.h:146-223) as the Sampler of the drop-in
    HybridLocallyOptimizedMSAC: the golden run of the reference's own headers, and a larger instance vs the oracle."""
    import json
    import os

    H = host()
    O = ol.oracle()
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_vectors.json")) as fh:
        h = json.load(fh)["hybrid"]
    xy0, p1 = np.array(h["xy0"]), np.array(h["xynn1"])
    w0, w1 = np.array(h["weights0"]), np.array(h["weights1"])
    for c in h["lomsac_biased"]:
        opts = H.HybridLORansacOptions(squared_inlier_thresholds_=[1e-4, 1e-4], data_type_weights_=[2.0, 0.5],
                                       random_seed_=c["seed"], final_least_squares_=bool(c["final_lsq"]),
                                       min_num_iterations_=100, max_num_iterations_=10000,
                                       max_num_iterations_per_solver_=1000)
        m, st = H.estimate_hybrid_line(opts, xy0, p1, [0.2, 0.8], sampling_weights=(w0, w1))
        assert st["num_iterations_total"] == c["num_iterations_total"]
        assert st["num_iterations_per_solver"] == c["num_iterations_per_solver"]
        assert st["best_num_inliers"] == c["best_num_inliers"] and st["best_solver_type"] == c["best_solver_type"]
        assert st["number_lo_iterations"] == c["number_lo_iterations"]
        np.testing.assert_allclose(st["best_model_score"], c["best_model_score"], rtol=1e-9)
        np.testing.assert_allclose(m, c["model"], rtol=1e-9, atol=1e-12)
        assert st["inlier_indices"][0].tolist() == c["inliers0"] and st["inlier_indices"][1].tolist() == c["inliers1"]

    n, seed = 20000, 11
    xy0, p1 = synth.hybrid_line_instance(900, n // 2, n // 2, n // 2, n // 2)
    rng = np.random.default_rng(5)
    w0, w1 = rng.uniform(0.0, 1.0, n), rng.uniform(0.0, 1.0, n)
    w0[rng.random(n) < 0.3] = 0.0
    o = ol.HybridOptions.default(sq_thr=[1e-4, 1e-4], weights=[2.0, 0.5], seed=seed, min_iters=100,
                                 max_iters=10000, max_iters_per_solver=1000)
    ost, om = ol.HybridStats(), np.zeros(3)
    a, b = np.zeros(n, np.int32), np.zeros(n, np.int32)
    O.hybrid_lomsac_line_biased(C.byref(o), n, ol.c64(xy0), w0, n, ol.c64(p1), w1, np.array([0.2, 0.8]), om,
                                C.byref(ost), a, b)
    opts = H.HybridLORansacOptions(squared_inlier_thresholds_=[1e-4, 1e-4], data_type_weights_=[2.0, 0.5],
                                   random_seed_=seed, min_num_iterations_=100, max_num_iterations_=10000,
                                   max_num_iterations_per_solver_=1000)
    for precision in (-1, 1):
        m, st = H.estimate_hybrid_line(opts, xy0, p1, [0.2, 0.8], sampling_weights=(w0, w1), precision=precision)
        assert st["num_iterations_total"] == ost.num_iterations_total
        assert st["num_iterations_per_solver"] == list(ost.num_iterations_per_solver)
        assert st["best_num_inliers"] == ost.best_num_inliers and st["number_lo_iterations"] == ost.number_lo_iterations
        np.testing.assert_allclose(st["best_model_score"], ost.best_model_score, rtol=1e-9)
        np.testing.assert_allclose(m, om, rtol=1e-9, atol=1e-12)
        assert st["inlier_indices"][0].tolist() == a[:ost.num_inliers_per_type[0]].tolist()
        assert st["inlier_indices"][1].tolist() == b[:ost.num_inliers_per_type[1]].tolist()


def test_hybrid_pose_estimation():
    """A hybrid POSE solver behind the drop-in HybridLocallyOptimizedMSAC (examples/hybrid_pose_estimator.h: 2D-3D and
    3D-3D matches, P3P and three-point solvers, a 12-double model, both residuals on the GPU as NVRTC source): the
    run must be the run of the reference's hybrid_ransac.h over the same solver (tests/test_oracle_golden.py pins the
    restated oracle to it) - iterations per solver, LO count, per-type inlier lists exact, model and score to 1e-9 -
    in f64 and in f32 mode, for three prior / seed / final-least-squares combinations."""
    H = host()
    O = ol.oracle()
    inst = synth.hybrid_pose_instance(31, 250, 250, 150, 150)
    f = inst["focal"]
    big = synth.hybrid_pose_instance(32, 4000, 6000, 1500, 2500)
    for data, seed, priors, final_lsq in ((inst, 1, (0.5, 0.5), False), (inst, 2, (0.9, 0.1), True),
                                          (inst, 3, (0.1, 0.9), False), (big, 4, (0.5, 0.5), False)):
        thr = (144.0, 0.05 ** 2)
        o = ol.HybridOptions.default(sq_thr=thr, weights=(1.0, 1.0), seed=seed, min_iters=100, max_iters=5000,
                                     max_iters_per_solver=5000, final_lsq=int(final_lsq))
        om, ost, a, b = ol.hybrid_lomsac_pose(O, o, f, f, data["points2D"], data["points3D"], data["cam1"],
                                              data["world1"], priors)
        opts = H.HybridLORansacOptions(squared_inlier_thresholds_=list(thr), data_type_weights_=[1.0, 1.0],
                                       random_seed_=seed, min_num_iterations_=100, max_num_iterations_=5000,
                                       max_num_iterations_per_solver_=5000, final_least_squares_=final_lsq)
        for precision in (-1, 1):
            m, st = H.estimate_hybrid_pose(opts, f, f, data["points2D"], data["points3D"], data["cam1"],
                                           data["world1"], priors, precision=precision)
            assert st["num_iterations_total"] == ost.num_iterations_total, (seed, precision)
            assert st["num_iterations_per_solver"] == list(ost.num_iterations_per_solver)
            assert st["best_num_inliers"] == ost.best_num_inliers and st["best_solver_type"] == ost.best_solver_type
            assert st["number_lo_iterations"] == ost.number_lo_iterations
            np.testing.assert_allclose(st["best_model_score"], ost.best_model_score, rtol=1e-9)
            np.testing.assert_allclose(m, om, rtol=1e-9, atol=1e-12)
            assert st["inlier_indices"][0].tolist() == a.tolist()
            assert st["inlier_indices"][1].tolist() == b.tolist()
        R = om[:9].reshape(3, 3).T
        assert np.abs(R - data["R_gt"]).max() < 5e-3 and np.abs(om[9:] - data["c_gt"]).max() < 5e-2


def test_hybrid_pose_degenerate_inputs():
    """The hybrid pose solver on inputs its minimal solvers cannot use: too few matches of either type, one data
    type empty, 3D-3D matches on one line (the triangle frames do not exist), all 2D-3D matches outliers. The
    drop-in layer must do what the reference's hybrid_ransac.h does with the same solver (the oracle) - including
    finding nothing and leaving the caller's model untouched."""
    H = host()
    O = ol.oracle()
    inst = synth.hybrid_pose_instance(41, 120, 120, 60, 60)
    f = inst["focal"]
    rng = np.random.default_rng(4)
    line = np.outer(np.linspace(1.0, 5.0, 80), [0.3, -0.2, 1.0])  # collinear in both frames
    cases = {
        "three 2D-3D, two 3D-3D": (inst["points2D"][:3], inst["points3D"][:3], inst["cam1"][:2], inst["world1"][:2]),
        "no 3D-3D matches": (inst["points2D"], inst["points3D"], np.zeros((0, 3)), np.zeros((0, 3))),
        "no 2D-3D matches": (np.zeros((0, 2)), np.zeros((0, 3)), inst["cam1"], inst["world1"]),
        "collinear 3D-3D, random 2D-3D": (rng.uniform(-300, 300, (50, 2)), rng.uniform(-5, 5, (50, 3)), line,
                                          line + [1.0, 2.0, 3.0]),
    }
    thr = (144.0, 0.05 ** 2)
    for name, (p2, p3, c1, w1) in cases.items():
        o = ol.HybridOptions.default(sq_thr=thr, weights=(1.0, 1.0), seed=3, min_iters=100, max_iters=400,
                                     max_iters_per_solver=400)
        om, ost, a, b = ol.hybrid_lomsac_pose(O, o, f, f, p2, p3, c1, w1, (0.5, 0.5))
        opts = H.HybridLORansacOptions(squared_inlier_thresholds_=list(thr), data_type_weights_=[1.0, 1.0],
                                       random_seed_=3, min_num_iterations_=100, max_num_iterations_=400,
                                       max_num_iterations_per_solver_=400)
        m, st = H.estimate_hybrid_pose(opts, f, f, p2, p3, c1, w1, (0.5, 0.5))
        assert st["num_iterations_total"] == ost.num_iterations_total, name
        assert st["num_iterations_per_solver"] == list(ost.num_iterations_per_solver), name
        assert st["best_num_inliers"] == ost.best_num_inliers, name
        np.testing.assert_allclose(m, om, rtol=1e-9, atol=1e-12, err_msg=name)
        assert st["inlier_indices"][0].tolist() == a.tolist() and st["inlier_indices"][1].tolist() == b.tolist(), name
        if ost.best_num_inliers == 0:
            assert m.tolist() == [0.0] * 12, name  # nothing found: the model is the caller's


def test_too_little_data_and_untouched_model():
    H = host()
    m, st = H.estimate_line(H.LORansacOptions(), np.array([[0.0, 1.0]]))  # 1 point < sample size 2
    assert st["best_num_inliers"] == 0 and not st["found"] and m.tolist() == [0, 0, 0]
    # coincident points: no minimal model exists -> nothing found, model left untouched
    m, st = H.estimate_line(H.LORansacOptions(max_num_iterations_=200), np.ones((50, 2)))
    assert st["best_num_inliers"] == 0 and not st["found"] and m.tolist() == [0, 0, 0]


def test_pyransaclib_localization():
    """pyransaclib.ransaclib_localization drop-in (examples/ransaclib_localization.cc:58-165)."""
    from ransaclib_b200 import pyransaclib

    assert pyransaclib.ransaclib_localization("q", 500.0, 500.0, np.zeros((3, 2)), np.zeros((3, 3)), 12.0, 10, 100,
                                              1000) == {"success": False}
    inst = synth.pose_instance(5000, 1000, 1000)
    f = np.float32(inst["focal"])
    r = pyransaclib.ransaclib_localization("q0", f, f, inst["points2D"], inst["points3D"], 12.0, 10, 100, 10000,
                                           random_seed=42)
    assert r["success"] and set(r) == {"success", "qvec", "tvec", "num_inliers", "inliers"}
    assert r["num_inliers"] == len(r["inliers"]) >= 990
    # pose close to ground truth: t = -R c
    R, c = inst["R_gt"], inst["c_gt"]
    np.testing.assert_allclose(r["tvec"], -R @ c, atol=0.02)
    w, x, y, z = r["qvec"]
    Rq = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                   [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                   [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
    np.testing.assert_allclose(Rq, R, atol=5e-3)
    # same options + seed through the oracle: identical inlier set
    kw = dict(sq_thr=float(np.float32(12.0)) ** 2, seed=42, final_lsq=1, lo_steps=10, min_iters=100, max_iters=10000,
              lo_start=60)
    om, ost, oinl = oracle_pose(inst["points2D"], inst["points3D"], float(f), **kw)
    assert r["inliers"] == oinl.tolist()


def test_copying_and_borrowing_constructors_agree(monkeypatch):
    """The example solvers either copy the caller's arrays (as the reference's estimators do) or, tagged
    gpu::BorrowData, keep pointers to them; the mirror borrows, RLB_HOST_COPY=1 selects the copies. Same runs."""
    H = host()
    xy = synth.line_instance(5, 2000, 1500)
    inst = synth.pose_instance(6, 800, 700)
    xy0, p1 = synth.hybrid_line_instance(7, 900, 600, 700, 500)
    lo = H.LORansacOptions(squared_inlier_threshold_=1e-4, random_seed_=3)
    po = H.LORansacOptions(squared_inlier_threshold_=144.0, random_seed_=4)
    ho = H.HybridLORansacOptions(squared_inlier_thresholds_=[1e-4, 1e-4], data_type_weights_=[2.0, 0.5], random_seed_=5,
                                 max_num_iterations_per_solver_=1000)

    def runs():
        f = inst["focal"]
        a = H.estimate_line(lo, xy)
        b = H.estimate_pose(po, f, f, inst["points2D"], inst["points3D"])
        c = H.estimate_hybrid_line(ho, xy0, p1, [0.2, 0.8])
        out = []
        for m, st in (a, b, c):
            idx = st["inlier_indices"]
            lists = idx if isinstance(idx, list) else [idx]
            out.append((np.asarray(m).tolist(), st["best_model_score"], [np.asarray(x).tolist() for x in lists]))
        return out

    borrowed = runs()
    monkeypatch.setenv("RLB_HOST_COPY", "1")
    copied = runs()
    assert borrowed == copied


def test_pose_f32_mode_far_from_the_origin_selects_the_same_models():
    """ADVICE (round 1): the f32 sweeps' "maybe better" guard must not miss a hypothesis that is better in f64 when
    pixel coordinates are large and the scene is far from the world origin. 8000 x 6000 images, focal 6000, depths
    50-400, the scene 2e4 units from the origin: the f32 run (tensor-core sweep for the large batches, FFMA2 for the
    small ones, every near-best hypothesis settled in f64) must be the f64 run, which must be the oracle's."""
    H = host()
    inst = synth.pose_instance(777, 3000, 3000, width=8000.0, height=6000.0, focal=6000.0, min_depth=50.0,
                               max_depth=400.0)
    shift = np.array([1.0e4, -2.0e4, 5.0e3])
    p2, p3 = inst["points2D"], inst["points3D"] + shift
    f = inst["focal"]
    for iters in (100, 6000):
        kw = dict(sq_thr=16.0, seed=11, min_iters=iters, max_iters=max(iters, 10000))
        om, ost, oinl = oracle_pose(p2, p3, f, **kw)
        opts = H.LORansacOptions(squared_inlier_threshold_=16.0, random_seed_=11, min_num_iterations_=iters,
                                 max_num_iterations_=max(iters, 10000))
        m64, st64 = H.estimate_pose(opts, f, f, p2, p3, precision=0)
        check_same(st64, m64, ost, om, oinl)
        m32, st32 = H.estimate_pose(opts, f, f, p2, p3, precision=1)
        check_same(st32, m32, ost, om, oinl)
