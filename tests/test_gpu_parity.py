"""GPU parity tests: the CUDA path, called through the C ABI, against the pinned oracle
on the same seeded inputs. Tolerances are BASELINE.json's: f64 scores <= 1e-9 relative,
counts / inlier sets / best id exact (points within 1e-9 relative of the threshold are
excused); f32 scores <= 1e-4 relative with the same selected model."""
import json
import os

import numpy as np
import pytest

import oracle_lib as ol
from ransaclib_b200 import synth

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
F64_RTOL = 1e-9
F32_RTOL = 1e-4


def rlb():
    import ransaclib_b200

    return ransaclib_b200


def assert_same_inliers(got, want, errs, thr):
    """Exact set equality, except indices whose squared error is within 1e-9 rel of thr."""
    got, want = set(map(int, got)), set(map(int, want))
    for i in got ^ want:
        assert abs(errs[i] - thr) <= 1e-9 * thr, f"index {i}: err {errs[i]} vs thr {thr}"


@pytest.fixture(scope="module")
def G():
    with open(os.path.join(HERE, "golden", "reference_vectors.json")) as fh:
        return json.load(fh)


@pytest.fixture(scope="module")
def pose_case():
    inst = synth.pose_instance(123, 1500, 1573)  # n = 3073: not a multiple of any tile size
    f = inst["focal"]
    O = ol.oracle()
    samples = ol.sample_uniform(O, 42, len(inst["points2D"]), 4, 3000)
    models, valid = ol.pose_minimal(O, f, f, 144.0, inst["points2D"], inst["points3D"], samples)
    return inst, f, samples, models, valid


# ------------------------------------------------------------------ golden fixtures
def test_golden_pose_scores(ctx, G):
    p = G["pose"]
    sc = ctx.pose_scorer(p["points2D"], p["points3D"], p["fx"], p["fy"])
    s, c, bid, bsc = sc.score_models(p["models"], p["sq_thr"], precision=rlb().RLB_F64)
    np.testing.assert_allclose(s, p["scores"], rtol=F64_RTOL)
    assert c.tolist() == p["counts"]
    assert bid == int(np.argmin(p["scores"])) and bsc == s[bid]
    s32, c32, bid32, _ = sc.score_models(p["models"], p["sq_thr"], precision=rlb().RLB_F32)
    np.testing.assert_allclose(s32, p["scores"], rtol=F32_RTOL)
    assert bid32 == bid
    for m, want in zip(p["models"][:3], p["inliers_first3"]):
        lists, counts, score = sc.get_inliers(m, p["sq_thr"], p["sq_thr"])
        assert lists[0].tolist() == want and counts[0] == len(want)
    np.testing.assert_allclose(sc.eval_errors(p["models"][1]), p["errors_model1"], rtol=1e-9)


def test_golden_line(ctx, G):
    l = G["line"]
    sc = ctx.line_scorer(l["xy"])
    m, v = sc.solve_line2pt(l["samples"])
    assert m.tolist() == l["models"] and v.tolist() == l["valid"]  # bit-exact (fmad off)
    s, c, bid, _ = sc.score_models(l["models"], l["sq_thr"])
    np.testing.assert_allclose(s, l["scores"], rtol=F64_RTOL)
    assert c.tolist() == l["counts"] and bid == l["best"]
    s32, _, bid32, _ = sc.score_models(l["models"], l["sq_thr"], precision=rlb().RLB_F32)
    np.testing.assert_allclose(s32, l["scores"], rtol=F32_RTOL)
    assert bid32 == l["best"]
    lists, counts, _ = sc.get_inliers(l["models"][l["best"]], l["sq_thr"])
    assert lists[0].tolist() == l["inliers_best"]


def test_golden_hybrid(ctx, G):
    h = G["hybrid"]
    sc = ctx.hybrid_line_scorer(h["xy0"], h["xynn1"])
    m0, v0 = sc.solve_line2pt(h["samples0"])
    m1, v1 = sc.solve_point_normal(h["samples1"])
    assert np.concatenate([m0, m1]).tolist() == h["models"]
    assert np.concatenate([v0, v1]).tolist() == h["valid"]
    s, c, bid, _ = sc.score_models(h["models"], h["sq_thr"], weights=h["weights"])
    np.testing.assert_allclose(s, h["scores"], rtol=F64_RTOL)
    assert c.tolist() == h["counts"] and bid == h["best"]
    s32, _, bid32, _ = sc.score_models(h["models"], h["sq_thr"], weights=h["weights"], precision=rlb().RLB_F32)
    np.testing.assert_allclose(s32, h["scores"], rtol=F32_RTOL)
    assert bid32 == h["best"]
    lists, counts, score = sc.get_inliers(h["models"][h["best"]], h["sq_thr"], h["sq_thr"], weights=h["weights"])
    assert lists[0].tolist() == h["inliers0"] and lists[1].tolist() == h["inliers1"]
    np.testing.assert_allclose(score, h["scores"][h["best"]], rtol=F64_RTOL)


# ------------------------------------------------------------- functor-level parity
def test_pose_errors_functor(ctx, pose_case):
    inst, f, _, models, valid = pose_case
    sc = ctx.pose_scorer(inst["points2D"], inst["points3D"], f, f)
    O = ol.oracle()
    n = len(inst["points2D"])
    for m in list(models[valid == 1][:5]) + [synth.pose_to_model(inst["R_gt"], inst["c_gt"])]:
        want = np.empty(n)
        O.pose_errors(ol.c64(m), f, f, n, inst["points2D"], inst["points3D"], want)
        got = sc.eval_errors(m)
        front = want < 1e300
        np.testing.assert_allclose(got[front], want[front], rtol=1e-9, atol=1e-12)
        assert (got[~front] > 1e300).all()  # behind-camera sentinel (DBL_MAX)
        got32 = sc.eval_errors(m, precision=rlb().RLB_F32)
        near = front & (want < 1e4)
        np.testing.assert_allclose(got32[near], want[near], rtol=2e-4, atol=1e-3)
        assert not (got32[~front] < 144.0).any()


def test_pose_sweep_f64_and_f32(ctx, pose_case):
    inst, f, _, models, valid = pose_case
    sc = ctx.pose_scorer(inst["points2D"], inst["points3D"], f, f)
    O = ol.oracle()
    want_s, want_c = ol.pose_score(O, f, f, inst["points2D"], inst["points3D"], models, 144.0)
    s, c, bid, bsc = sc.score_models(models, 144.0, valid=valid)
    ok = valid == 1
    np.testing.assert_allclose(s[ok], want_s[ok], rtol=F64_RTOL)
    assert (c[ok] == want_c[ok]).all()
    assert (s[~ok] == np.finfo(np.float64).max).all() and (c[~ok] == 0).all()
    masked = np.where(ok, want_s, np.inf)
    assert bid == int(np.argmin(masked)) and bsc == s[bid]
    s32, c32, bid32, _ = sc.score_models(models, 144.0, valid=valid, precision=rlb().RLB_F32)
    np.testing.assert_allclose(s32[ok], want_s[ok], rtol=F32_RTOL)
    assert bid32 == bid
    assert np.abs(c32[ok] - want_c[ok]).max() <= 2  # only points numerically on the threshold


def test_pose_sweep_without_counts(ctx, pose_case):
    """counts == NULL (what ScoreModel computes, ransac.h:300-312): same scores and winner as with counts."""
    inst, f, _, models, valid = pose_case
    sc = ctx.pose_scorer(inst["points2D"], inst["points3D"], f, f)
    O = ol.oracle()
    want_s, _ = ol.pose_score(O, f, f, inst["points2D"], inst["points3D"], models, 144.0)
    ok = valid == 1
    for prec, tol in ((rlb().RLB_F64, F64_RTOL), (rlb().RLB_F32, F32_RTOL)):
        s, c, bid, bsc = sc.score_models(models, 144.0, valid=valid, precision=prec, want_counts=False)
        assert c is None
        np.testing.assert_allclose(s[ok], want_s[ok], rtol=tol)
        s2, _, bid2, bsc2 = sc.score_models(models, 144.0, valid=valid, precision=prec)
        np.testing.assert_allclose(s[ok], s2[ok], rtol=1e-6 if prec == rlb().RLB_F32 else 0)
        assert bid == bid2 == int(np.argmin(np.where(ok, want_s, np.inf)))


def test_pose_inliers_ordered_and_fused_score(ctx, pose_case):
    inst, f, _, models, valid = pose_case
    sc = ctx.pose_scorer(inst["points2D"], inst["points3D"], f, f)
    O = ol.oracle()
    n = len(inst["points2D"])
    best = models[valid == 1][0]
    for m in [best, synth.pose_to_model(inst["R_gt"], inst["c_gt"])]:
        for thr_inl, thr_sc in [(144.0, 144.0), (144.0 * 2 ** 0.5, 144.0), (1.0, 144.0)]:
            lists, counts, score = sc.get_inliers(m, thr_inl, thr_sc)
            want = ol.pose_inliers(O, f, f, inst["points2D"], inst["points3D"], m, thr_inl)
            errs = np.empty(n)
            O.pose_errors(ol.c64(m), f, f, n, inst["points2D"], inst["points3D"], errs)
            assert_same_inliers(lists[0], want, errs, thr_inl)
            assert (np.diff(lists[0]) > 0).all()  # ascending (ransac.h:327-333)
            assert counts[0] == len(lists[0])
            ws, _ = ol.pose_score(O, f, f, inst["points2D"], inst["points3D"], m[None], thr_sc)
            np.testing.assert_allclose(score, ws[0], rtol=F64_RTOL)
    # count-only mode
    _, counts, _ = sc.get_inliers(best, 144.0, want_indices=False)
    assert counts[0] == len(ol.pose_inliers(O, f, f, inst["points2D"], inst["points3D"], best, 144.0))


def test_p3p_batch_bit_exact(ctx, pose_case):
    inst, f, samples, models, valid = pose_case
    sc = ctx.pose_scorer(inst["points2D"], inst["points3D"], f, f)
    O = ol.oracle()
    rays = np.empty((len(inst["points2D"]), 3))
    O.pixels_to_rays(f, f, len(rays), inst["points2D"], rays)
    assert sc.rays().tolist() == rays.tolist()
    gm, gv = sc.solve_p3p(samples, 144.0)
    assert gv.tolist() == valid.tolist()
    assert gm.tolist() == models.tolist()  # no FMA contraction on either side: identical bits


def test_hypothesize_and_score(ctx, pose_case):
    inst, f, samples, models, valid = pose_case
    sc = ctx.pose_scorer(inst["points2D"], inst["points3D"], f, f)
    O = ol.oracle()
    want_s, want_c = ol.pose_score(O, f, f, inst["points2D"], inst["points3D"], models, 144.0)
    masked = np.where(valid == 1, want_s, np.inf)
    for prec, rtol in [(rlb().RLB_F64, F64_RTOL), (rlb().RLB_F32, F32_RTOL)]:
        r = sc.hypothesize_and_score(samples, 144.0, precision=prec)
        assert r["valid"].tolist() == valid.tolist()
        assert r["models"].tolist() == models.tolist()
        ok = valid == 1
        np.testing.assert_allclose(r["scores"][ok], want_s[ok], rtol=rtol)
        assert r["best_id"] == int(np.argmin(masked))
        assert r["best_model"].tolist() == models[r["best_id"]].tolist()


# --------------------------------------------------------------------- edge cases
def test_ties_invalid_nan_and_tiny_inputs(ctx):
    R = rlb()
    inst = synth.pose_instance(5, 4, 1)  # n = 5 (< one vector group, < one stage)
    f = inst["focal"]
    sc = ctx.pose_scorer(inst["points2D"], inst["points3D"], f, f)
    gt = synth.pose_to_model(inst["R_gt"], inst["c_gt"])
    bad = gt.copy()
    bad[3] = np.nan
    models = np.stack([bad, gt, gt, gt])  # NaN model first, then a three-way tie
    s, c, bid, bsc = sc.score_models(models, 144.0)
    assert np.isnan(s[0]) and c[0] == 0  # NaN propagates and never wins (ransac.h:308)
    assert bid == 1 and s[1] == s[2] == s[3]  # lowest id wins ties (ransac.h:288)
    s, c, bid, _ = sc.score_models(models, 144.0, valid=np.array([1, 0, 0, 1], np.uint8))
    assert bid == 3
    s, c, bid, bsc = sc.score_models(models, 144.0, valid=np.zeros(4, np.uint8))
    assert bid == -1 and bsc == np.finfo(np.float64).max
    O = ol.oracle()
    ws, wc = ol.pose_score(O, f, f, inst["points2D"], inst["points3D"], gt[None], 144.0)
    s1, c1, b1, _ = sc.score_models(gt[None], 144.0)  # K = 1
    np.testing.assert_allclose(s1, ws, rtol=F64_RTOL)
    assert c1.tolist() == wc.tolist() and b1 == 0
    with pytest.raises(R.RlbError):
        sc.solve_p3p(np.array([[0, 1, 2, 5]], np.int32), 144.0)  # index out of range


def test_empty_and_ragged_hybrid(ctx):
    xy0, p1 = synth.hybrid_line_instance(3, 700, 500, 10, 7)  # 1200 / 17 points: ragged segments
    sc = ctx.hybrid_line_scorer(xy0, p1)
    O = ol.oracle()
    smp = ol.sample_uniform(O, 1, len(xy0), 2, 64)
    models, valid = sc.solve_line2pt(smp)
    thr, w = np.array([1e-4, 3e-4]), np.array([2.0, 0.5])
    ws, wc = np.empty(64), np.empty(64, np.int32)
    O.hybrid_line_score_models(len(xy0), ol.c64(xy0), len(p1), ol.c64(p1), 64, ol.c64(models), thr, w, ws, wc)
    s, c, bid, _ = sc.score_models(models, thr, weights=w)
    np.testing.assert_allclose(s, ws, rtol=F64_RTOL)
    assert c.tolist() == wc.tolist() and bid == int(np.argmin(ws))
    # second data type empty
    sc2 = ctx.hybrid_line_scorer(xy0, np.zeros((0, 4)))
    s2, c2, _, _ = sc2.score_models(models, thr, weights=w)
    ws2, wc2 = np.empty(64), np.empty(64, np.int32)
    O.line_score_models(len(xy0), ol.c64(xy0), 64, ol.c64(models), thr[0], ws2, wc2)
    np.testing.assert_allclose(s2, 2.0 * ws2, rtol=F64_RTOL)
    lists, counts, _ = sc2.get_inliers(models[int(np.argmin(ws))], thr)
    assert counts.tolist() == [len(lists[0]), 0]


def test_small_batches_over_many_points(ctx):
    """The sweep plan of a batch that cannot fill the device with 64 point-axis splits (the drivers' rounds over a
    large data set): one hypothesis per thread for the line kinds, up to 256 splits - scores and counts against the
    oracle at batch sizes around the tile edges, both arithmetic classes, line / hybrid / pose."""
    O = ol.oracle()
    n = 150000
    xy = synth.line_instance(31, n // 2, n - n // 2)
    sc = ctx.line_scorer(xy)
    models, _ = sc.solve_line2pt(ol.sample_uniform(O, 5, n, 2, 600))
    for k in (1, 15, 128, 129, 400, 513, 600):
        ws, wc = np.empty(k), np.empty(k, np.int32)
        O.line_score_models(n, ol.c64(xy), k, ol.c64(models[:k]), 1e-4, ws, wc)
        s, c, bid, _ = sc.score_models(models[:k], 1e-4)
        variant, splits = ctx.last_sweep_plan()
        assert splits > 64 and (variant == 10 or k > 128), (k, variant, splits)
        np.testing.assert_allclose(s, ws, rtol=F64_RTOL)
        assert c.tolist() == wc.tolist() and bid == int(np.argmin(ws))
        s32, c32, _, _ = sc.score_models(models[:k], 1e-4, precision=rlb().RLB_F32)
        np.testing.assert_allclose(s32, ws, rtol=F32_RTOL)
        assert np.abs(c32 - wc).max() <= max(2, int(2e-5 * n))  # points within f32 rounding of the threshold

    xy0, p1 = synth.hybrid_line_instance(8, 40000, 30000, 50000, 45000)
    hs = ctx.hybrid_line_scorer(xy0, p1)
    hm, _ = hs.solve_line2pt(ol.sample_uniform(O, 6, len(xy0), 2, 100))
    thr, w = np.array([1e-4, 3e-4]), np.array([2.0, 0.5])
    ws, wc = np.empty(100), np.empty(100, np.int32)
    O.hybrid_line_score_models(len(xy0), ol.c64(xy0), len(p1), ol.c64(p1), 100, ol.c64(hm), thr, w, ws, wc)
    s, c, bid, _ = hs.score_models(hm, thr, weights=w)
    assert ctx.last_sweep_plan()[1] > 64
    np.testing.assert_allclose(s, ws, rtol=F64_RTOL)
    assert c.tolist() == wc.tolist() and bid == int(np.argmin(ws))

    inst = synth.pose_instance(77, 60000, 60000)
    f = inst["focal"]
    ps = ctx.pose_scorer(inst["points2D"], inst["points3D"], f, f)
    pm = synth.perturbed_poses(3, inst["R_gt"], inst["c_gt"], 96, rot_sigma=0.003, trans_sigma=0.01)
    ws, wc = ol.pose_score(O, f, f, inst["points2D"], inst["points3D"], pm, 144.0)
    s, c, bid, _ = ps.score_models(pm, 144.0)
    assert ctx.last_sweep_plan()[1] > 64
    np.testing.assert_allclose(s, ws, rtol=F64_RTOL)
    assert c.tolist() == wc.tolist() and bid == int(np.argmin(ws))
    s32, _, _, _ = ps.score_models(pm, 144.0, precision=rlb().RLB_F32)
    np.testing.assert_allclose(s32, ws, rtol=F32_RTOL)


def test_host_p3p_is_the_device_p3p(ctx, pose_case):
    """examples/p3p.h compiles for both sides: the estimator's host MinimalSolver (NonMinimalSolver calls it once per
    LO step, no launch) returns the bits of the device solver, valid and rejected samples alike."""
    from ransaclib_b200 import host as H

    inst, f, samples, models, valid = pose_case
    hm, hv = H.pose_minimal_solver(f, f, 144.0, inst["points2D"], inst["points3D"], samples)
    assert hv.tolist() == valid.tolist()
    assert hm.tolist() == models.tolist()
    more = ol.sample_uniform(ol.oracle(), 99, len(inst["points2D"]), 4, 3000)
    sc = ctx.pose_scorer(inst["points2D"], inst["points3D"], f, f)
    gm, gv = sc.solve_p3p(more, 144.0)
    hm, hv = H.pose_minimal_solver(f, f, 144.0, inst["points2D"], inst["points3D"], more)
    assert hv.tolist() == gv.tolist() and hm.tolist() == gm.tolist()


def test_inlier_list_view(ctx):
    """rlb_get_inliers_view: the list read in place from the context's pinned copy is the list rlb_get_inliers copies
    out - per type for a hybrid scorer, empty lists, and past the zero-copy limit (262,144 points) where the copy
    is fetched by DMA."""
    xy0, p1 = synth.hybrid_line_instance(5, 900, 600, 400, 300)
    sc = ctx.hybrid_line_scorer(xy0, p1)
    smp = ol.sample_uniform(ol.oracle(), 2, len(xy0), 2, 8)
    models, _ = sc.solve_line2pt(smp)
    thr, w = np.array([1e-4, 3e-4]), np.array([2.0, 0.5])
    for m in models[:4]:
        for t in (thr, 50.0 * thr, 0.0 * thr, -thr):
            lists, counts, score = sc.get_inliers(m, t, thr, w)
            vl, vc, vs = sc.get_inliers_view(m, t, thr, w)
            assert vc.tolist() == counts.tolist() and vs == score
            assert [x.tolist() for x in vl] == [x.tolist() for x in lists]
            if (t <= 0).all():
                assert counts.tolist() == [0, 0]
    n = 300000
    xy = synth.line_instance(77, n // 2, n - n // 2)
    big = ctx.line_scorer(xy)
    bm, _ = big.solve_line2pt(ol.sample_uniform(ol.oracle(), 3, n, 2, 4))
    for m in bm:
        lists, counts, score = big.get_inliers(m, 1e-3, 1e-4)
        vl, vc, vs = big.get_inliers_view(m, 1e-3, 1e-4)
        assert vc.tolist() == counts.tolist() and vs == score and vl[0].tolist() == lists[0].tolist()
    _, vc, _ = big.get_inliers_view(bm[0], 0.0)
    assert vc.tolist() == [0]


def test_far_from_origin_f32_recentring(ctx):
    """Scene 1e5 units from the origin: the f32 path must still meet 1e-4 (points are
    recentred in f64 before the f32 conversion)."""
    inst = synth.pose_instance(9, 1000, 1000)
    shift = np.array([1.2e5, -3.4e5, 5.6e4])
    p3 = inst["points3D"] + shift
    f = inst["focal"]
    sc = ctx.pose_scorer(inst["points2D"], p3, f, f)
    models = synth.perturbed_poses(2, inst["R_gt"], inst["c_gt"] + shift, 64, rot_sigma=0.003, trans_sigma=0.01)
    O = ol.oracle()
    ws, wc = ol.pose_score(O, f, f, inst["points2D"], p3, models, 144.0)
    s, c, bid, _ = sc.score_models(models, 144.0)
    np.testing.assert_allclose(s, ws, rtol=1e-7)  # f64 input coordinates only carry ~1e-11 rel here
    s32, _, bid32, _ = sc.score_models(models, 144.0, precision=rlb().RLB_F32)
    np.testing.assert_allclose(s32, ws, rtol=F32_RTOL)
    assert bid32 == int(np.argmin(ws))


@pytest.mark.parametrize("n", [5, 31, 32, 33, 700, 4097])
def test_tensor_core_sweep_edges(ctx, n):
    """The tensor-core sweep (f32 pose, K >= 512) on point counts around its 32-point tile, with hypotheses
    that put every point behind the camera, and against the FFMA2 sweep (K < 512) on the same models."""
    n_in = max(3, n // 2)
    inst = synth.pose_instance(77 + n, n_in, n - n_in)
    f = inst["focal"]
    p2, p3 = inst["points2D"], inst["points3D"]
    sc = ctx.pose_scorer(p2, p3, f, f)
    models = synth.perturbed_poses(3, inst["R_gt"], inst["c_gt"], 640, rot_sigma=0.01, trans_sigma=0.05)
    behind = models[:8].copy().reshape(8, 4, 3)  # column-major [R | c]: behind[m, j] is column j
    behind[:, :3, 1:] *= -1.0  # R' = diag(1, -1, -1) R, half a turn about x: every depth changes sign
    models[100:108] = behind.reshape(8, 12)
    O = ol.oracle()
    ws, wc = ol.pose_score(O, f, f, p2, p3, models, 144.0)
    assert (wc[100:108] == 0).all() and np.allclose(ws[100:108], 144.0 * n)
    for k in (640, 513, 512):
        s, c, bid, _ = sc.score_models(models[:k], 144.0, precision=rlb().RLB_F32)
        assert ctx.last_sweep_plan()[0] == 9
        np.testing.assert_allclose(s, ws[:k], rtol=F32_RTOL)
        assert np.abs(c - wc[:k]).max() <= 2 and (c[100:108] == 0).all()
        assert bid == int(np.argmin(ws[:k]))
        s_nc, c_nc, bid_nc, _ = sc.score_models(models[:k], 144.0, precision=rlb().RLB_F32, want_counts=False)
        np.testing.assert_allclose(s_nc, s, rtol=1e-6)
        assert c_nc is None and bid_nc == bid
    s_f, c_f, _, _ = sc.score_models(models[:500], 144.0, precision=rlb().RLB_F32)
    assert ctx.last_sweep_plan()[0] != 9  # below 512 hypotheses the FFMA2 sweep runs
    np.testing.assert_allclose(s_f, s[:500], rtol=2e-5)
    assert np.abs(c_f - c[:500]).max() <= 2


@pytest.mark.parametrize("n", [5, 31, 32, 33, 700, 4097, 70_000])
def test_sorted_two_speed_sweep(ctx, n):
    """RLB_F32_SORTED (sweep variant 12, an opt-in experiment: RLB_TC_SORTED=1 lets RLB_F32 pick it): -u / -v
    inside the tensor-core contraction, a reciprocal-free filter on every pair, the exact f32 evaluation only where
    a warp has a flagged hypothesis, hypotheses swept in the order of a prepass. Against the oracle (f32 tolerance,
    same winner), against the unsorted tensor-core sweep, with and without counts; good, bad, behind-the-camera,
    invalid and non-finite hypotheses mixed so that warps hold both kinds; and a hypothesis' inlier decisions must
    not depend on what else is in the batch."""
    n_in = max(3, n // 2)
    inst = synth.pose_instance(277 + n, n_in, n - n_in)
    f = inst["focal"]
    p2, p3 = inst["points2D"], inst["points3D"]
    sc = ctx.pose_scorer(p2, p3, f, f)
    k = 1300
    rng = np.random.default_rng(n)
    good = synth.perturbed_poses(5, inst["R_gt"], inst["c_gt"], k, rot_sigma=0.002, trans_sigma=0.01)
    bad = synth.perturbed_poses(6, inst["R_gt"], inst["c_gt"], k, rot_sigma=0.6, trans_sigma=3.0)
    models = np.where((rng.random(k) < 0.15)[:, None], good, bad)
    behind = good[:8].copy().reshape(8, 4, 3)  # half a turn about x: every depth changes sign
    behind[:, :3, 1:] *= -1.0
    models[100:108] = behind.reshape(8, 12)
    models[200, 4] = np.nan
    models[201, 10] = np.inf
    valid = np.ones(k, dtype=np.uint8)
    valid[300:310] = 0
    O = ol.oracle()
    ws, wc = ol.pose_score(O, f, f, p2, p3, models, 144.0)
    ok = np.ones(k, dtype=bool)
    ok[[200, 201]] = False
    ok[300:310] = False
    m = rlb()
    s, c, bid, _ = sc.score_models(models, 144.0, valid=valid, precision=m.RLB_F32_SORTED)
    assert ctx.last_sweep_plan()[0] == 12
    np.testing.assert_allclose(s[ok], ws[ok], rtol=F32_RTOL)
    assert np.abs(c[ok] - wc[ok]).max() <= 2 and (c[100:108] == 0).all()
    assert np.isnan(s[200]) and (s[300:310] == np.finfo(np.float64).max).all()
    wsm = np.where(ok, ws, np.inf)
    assert bid == int(np.argmin(wsm))
    s9, c9, bid9, _ = sc.score_models(models, 144.0, valid=valid, precision=m.RLB_F32_UNSORTED)
    assert ctx.last_sweep_plan()[0] == 9
    np.testing.assert_allclose(s[ok], s9[ok], rtol=2e-5)
    assert np.abs(c[ok] - c9[ok]).max() <= 2 and bid9 == bid
    s_nc, c_nc, bid_nc, _ = sc.score_models(models, 144.0, valid=valid, precision=m.RLB_F32_SORTED, want_counts=False)
    assert c_nc is None and bid_nc == bid
    np.testing.assert_array_equal(s_nc[ok], s[ok])
    # another batch composition (other neighbours in the warp, another sweep order): the same inlier decisions;
    # the sums agree to f32-flush rounding only (another K gives another split plan, as for every sweep here)
    sub = np.concatenate([np.arange(0, k, 3), np.arange(1, k, 7)])
    s_sub, c_sub, _, _ = sc.score_models(models[sub], 144.0, valid=valid[sub], precision=m.RLB_F32_SORTED)
    np.testing.assert_allclose(s_sub[ok[sub]], s[sub][ok[sub]], rtol=1e-7)
    np.testing.assert_array_equal(c_sub[ok[sub]], c[sub][ok[sub]])


@pytest.mark.parametrize("n", [5, 33, 700, 4097, 70_000])
def test_filtered_f64_sweep_equals_the_plain_f64_sweep(ctx, n):
    """RLB_F64_FILTERED (what RLB_F64 runs for pose batches of K x N >= 2^28): f32 classification on the tensor
    cores, f64 only for the pairs that could be inliers (sweep variant 11). Counts must be those of the plain f64 sweep (RLB_F64_PLAIN) exactly, scores to
    1e-12 (fixed-point sums), both within 1e-9 of the oracle; good, bad, behind-the-camera, invalid, non-finite
    and grazing hypotheses (camera centre inside the point cloud: depths around zero)."""
    n_in = max(3, n // 2)
    inst = synth.pose_instance(177 + n, n_in, n - n_in)
    f = inst["focal"]
    p2, p3 = inst["points2D"], inst["points3D"]
    sc = ctx.pose_scorer(p2, p3, f, f)
    k = 700
    models = synth.perturbed_poses(5, inst["R_gt"], inst["c_gt"], k, rot_sigma=0.01, trans_sigma=0.05)
    rng = np.random.default_rng(n)
    behind = models[:8].copy().reshape(8, 4, 3)
    behind[:, :3, 1:] *= -1.0
    models[100:108] = behind.reshape(8, 12)
    # camera centres ON data points / inside the cloud: zero and tiny depths, points on both sides
    for i in range(200, 232):
        m = models[i].copy()
        m[9:12] = p3[rng.integers(0, n)] + (rng.normal(size=3) * 1e-3 if i % 2 else 0.0)
        models[i] = m
    for i in range(300, 364):  # arbitrary rotations about the true centre and random centres
        R = synth.random_rotation(rng)
        c = inst["c_gt"] + rng.normal(size=3) * (0.0 if i < 332 else 3.0)
        models[i] = synth.pose_to_model(R, c)
    valid = np.ones(k, dtype=np.uint8)
    valid[50:60] = 0
    models[60, 3] = np.nan
    models[61, 10] = np.inf
    s_p, c_p, bid_p, _ = sc.score_models(models, 144.0, valid=valid, precision=rlb().RLB_F64_PLAIN)
    assert ctx.last_sweep_plan()[0] != 11
    s_f, c_f, bid_f, _ = sc.score_models(models, 144.0, valid=valid, precision=rlb().RLB_F64_FILTERED)
    assert ctx.last_sweep_plan()[0] == 11
    assert c_f.tolist() == c_p.tolist()
    fin = np.isfinite(s_p) & (s_p < 1e300)
    assert (np.isnan(s_f) == np.isnan(s_p)).all() and ((s_f > 1e300) == (s_p > 1e300)).all()
    np.testing.assert_allclose(s_f[fin], s_p[fin], rtol=1e-12)
    assert bid_f == bid_p
    O = ol.oracle()
    pick = np.r_[0:40, 100:108, 200:232, 300:364]
    ws, wc = ol.pose_score(O, f, f, p2, p3, models[pick], 144.0)
    ok = np.isfinite(ws)
    np.testing.assert_allclose(s_f[pick][ok], ws[ok], rtol=F64_RTOL)
    assert c_f[pick][ok].tolist() == wc[ok].tolist()
    # other thresholds (tiny: the band is relatively wide; huge: nearly everything is evaluated in f64)
    for thr in (0.25, 1e4):
        a = sc.score_models(models, thr, valid=valid, precision=rlb().RLB_F64_PLAIN)
        b = sc.score_models(models, thr, valid=valid, precision=rlb().RLB_F64_FILTERED)
        assert a[1].tolist() == b[1].tolist() and a[2] == b[2]
        fin = np.isfinite(a[0]) & (a[0] < 1e300)
        np.testing.assert_allclose(b[0][fin], a[0][fin], rtol=1e-12)
    # bit-reproducible
    again = sc.score_models(models, 144.0, valid=valid, precision=rlb().RLB_F64_FILTERED)
    assert np.array_equal(again[0], s_f, equal_nan=True)


def test_filtered_f64_sweep_far_from_the_origin(ctx):
    """Large pixel coordinates and a scene far from the world origin (ADVICE: the f32 guard must come from an
    error bound, not a constant): the per-hypothesis band of the f32 classification still covers every inlier."""
    n = 20_000
    inst = synth.pose_instance(4242, n // 2, n // 2, width=8000.0, height=6000.0, focal=6000.0,
                               min_depth=50.0, max_depth=400.0)
    f = inst["focal"]
    shift = np.array([1.0e4, -2.0e4, 5.0e3])
    p2, p3 = inst["points2D"], inst["points3D"] + shift
    sc = ctx.pose_scorer(p2, p3, f, f)
    models = synth.perturbed_poses(9, inst["R_gt"], inst["c_gt"] + shift, 1024, rot_sigma=0.002, trans_sigma=0.5)
    for thr in (4.0, 144.0):
        a = sc.score_models(models, thr, precision=rlb().RLB_F64_PLAIN)
        b = sc.score_models(models, thr, precision=rlb().RLB_F64_FILTERED)
        assert ctx.last_sweep_plan()[0] == 11
        assert a[1].tolist() == b[1].tolist() and a[2] == b[2]
        np.testing.assert_allclose(b[0], a[0], rtol=1e-12)
    O = ol.oracle()
    ws, wc = ol.pose_score(O, f, f, p2, p3, models[:64], 144.0)
    np.testing.assert_allclose(b[0][:64], ws, rtol=F64_RTOL)
    assert b[1][:64].tolist() == wc.tolist()


def test_f64_picks_the_filtered_sweep_by_size(ctx):
    """RLB_F64 = the plain f64 sweep for small batches (one launch), the filtered one from K x N >= 2^28."""
    inst = synth.pose_instance(91, 150_000, 150_000)
    f = inst["focal"]
    sc = ctx.pose_scorer(inst["points2D"], inst["points3D"], f, f)
    models = synth.perturbed_poses(2, inst["R_gt"], inst["c_gt"], 1024)
    a = sc.score_models(models[:512], 144.0)  # 512 x 300k < 2^28
    assert ctx.last_sweep_plan()[0] != 11
    b = sc.score_models(models, 144.0)  # 1024 x 300k >= 2^28
    assert ctx.last_sweep_plan()[0] == 11
    assert b[1][:512].tolist() == a[1].tolist()
    np.testing.assert_allclose(b[0][:512], a[0], rtol=1e-12)


def test_determinism(ctx, pose_case):
    inst, f, _, models, valid = pose_case
    sc = ctx.pose_scorer(inst["points2D"], inst["points3D"], f, f)
    a = sc.score_models(models, 144.0, precision=rlb().RLB_F32)
    b = sc.score_models(models, 144.0, precision=rlb().RLB_F32)
    assert a[0].tolist() == b[0].tolist() and a[1].tolist() == b[1].tolist()
    l1 = sc.get_inliers(models[valid == 1][0], 144.0, 144.0)
    l2 = sc.get_inliers(models[valid == 1][0], 144.0, 144.0)
    assert l1[0][0].tolist() == l2[0][0].tolist() and l1[2] == l2[2]


# ------------------------------------------- full-size, size-independent properties
def test_full_size_properties(ctx):
    """BASELINE config sizes (N = 1M points) are too slow for the scalar oracle on all K;
    check (a) a sampled subset of hypotheses against the oracle, (b) linearity: scoring the
    two halves of the point set separately sums to the whole, (c) idempotence."""
    n, k = 1_000_000, 4096
    inst = synth.pose_instance(9000, n // 2, n // 2)
    f = inst["focal"]
    sc = ctx.pose_scorer(inst["points2D"], inst["points3D"], f, f)
    models = synth.perturbed_poses(1, inst["R_gt"], inst["c_gt"], k)
    s, c, bid, bsc = sc.score_models(models, 144.0, precision=rlb().RLB_F32)
    s64, c64_, bid64, _ = sc.score_models(models, 144.0)
    np.testing.assert_allclose(s, s64, rtol=F32_RTOL)
    assert bid == bid64
    O = ol.oracle()
    pick = np.array([0, 1, 7, bid, k - 1])
    ws, wc = ol.pose_score(O, f, f, inst["points2D"], inst["points3D"], models[pick], 144.0)
    np.testing.assert_allclose(s64[pick], ws, rtol=F64_RTOL)
    assert c64_[pick].tolist() == wc.tolist()
    h = n // 2 + 137
    sa = ctx.pose_scorer(inst["points2D"][:h], inst["points3D"][:h], f, f)
    sb = ctx.pose_scorer(inst["points2D"][h:], inst["points3D"][h:], f, f)
    a = sa.score_models(models[:256], 144.0)
    b = sb.score_models(models[:256], 144.0)
    np.testing.assert_allclose(a[0] + b[0], s64[:256], rtol=1e-9)
    assert (a[1] + b[1]).tolist() == c64_[:256].tolist()
    lists, counts, score = sc.get_inliers(models[bid64], 144.0, 144.0)
    assert counts[0] == c64_[bid64] and (np.diff(lists[0]) > 0).all()
    np.testing.assert_allclose(score, s64[bid64], rtol=F64_RTOL)


def test_sharded_round_emulated_ranks(ctx, pose_case):
    """Hypothesis sharding with fewer GPUs than ranks: the ranks' shards run one
    after the other on this GPU, their records are stacked as the all_gather would, and the winner
    must be the single-GPU winner (same global id, same model) for every world size."""
    import torch

    from ransaclib_b200 import sharding

    inst, f, samples, models, valid = pose_case
    sc = ctx.pose_scorer(inst["points2D"], inst["points3D"], f, f)
    ref = sc.hypothesize_and_score(samples, 144.0)
    dev = torch.device("cuda", 0)
    k_total = len(samples)
    for world in (1, 2, 3, 8):
        recs = torch.zeros((world, 14), dtype=torch.float64, device=dev)
        for rank in range(world):
            lo, hi = sharding.hypothesis_range(k_total, rank, world)
            k = hi - lo
            d_s = torch.from_numpy(np.ascontiguousarray(samples[lo:hi])).to(dev)
            d_m = torch.empty((k, 12), dtype=torch.float64, device=dev)
            d_v = torch.empty(k, dtype=torch.uint8, device=dev)
            d_sc = torch.empty(k, dtype=torch.float64, device=dev)
            d_c = torch.empty(k, dtype=torch.int32, device=dev)
            torch.cuda.synchronize()
            sc.sharded_round_dev(d_s.data_ptr(), k, 144.0, None, lo, d_m.data_ptr(), d_v.data_ptr(),
                                 d_sc.data_ptr(), d_c.data_ptr(), recs[rank].data_ptr())
            ctx.synchronize()
        win = torch.zeros(14, dtype=torch.float64, device=dev)
        ctx.select_winner_dev(recs.data_ptr(), world, 14, win.data_ptr())
        ctx.synchronize()
        w = win.cpu().numpy()
        # (a shard below 512 hypotheses runs the plain f64 sweep, the whole batch the filtered one: the same
        # pairs summed in another order)
        assert int(w[1]) == ref["best_id"] and abs(w[0] - ref["best_score"]) <= 1e-12 * ref["best_score"]
        assert w[2:].tolist() == ref["best_model"].tolist()
        # the torch-side selection used under gloo/NCCL agrees with the kernel
        row, ok = sharding.select_winner(recs)
        assert bool(ok) and int(recs[row, 1].item()) == ref["best_id"]


def test_zero_depth_falls_through_like_the_reference(ctx):
    """p.z == 0 is NOT behind the camera (`p.z < 0` only): the reference divides by
    zero, so the error is +inf (px != 0: contributes thr, not an inlier) or NaN (px == py == 0: 0/0,
    which poisons the score through std::min). The f64 kernels replace the IEEE division by a seeded
    reciprocal and must keep exactly this behaviour. One-point data sets make the recentred depth an
    exact zero."""
    O = ol.oracle()
    f = 100.0
    R = np.eye(3)
    for X, want_nan in (((1.0, 2.0, 5.0), False), ((0.0, 0.0, 5.0), True)):
        p2 = np.array([[3.0, 4.0]])
        p3 = np.array([X])
        model = synth.pose_to_model(R, np.array([0.0, 0.0, 5.0]))  # depth of the point: exactly 0
        want_s, want_c = ol.pose_score(O, f, f, p2, p3, model[None, :], 144.0)
        sc = ctx.pose_scorer(p2, p3, f, f)
        s, c, bid, bsc = sc.score_models(model[None, :], 144.0)
        err = sc.eval_errors(model)
        lists, counts, score = sc.get_inliers(model, 144.0, 144.0)
        assert counts[0] == 0 and want_c[0] == 0 and c[0] == 0
        if want_nan:
            assert np.isnan(want_s[0]) and np.isnan(s[0]) and np.isnan(err[0]) and np.isnan(score)
            assert bid == -1  # a NaN score never wins (ransac.h:195 is a strict <)
        else:
            assert want_s[0] == 144.0 and s[0] == 144.0 and score == 144.0
            assert np.isinf(err[0]) and err[0] > 0
        sc.close()


@pytest.mark.parametrize("n", [607_000, 1_300_000, 2_600_000, 4_850_000])
def test_inlier_pass_one_wave_of_tile_ranges(ctx, n):
    """The K = 1 pass as one wave (inlier_range_kernel): 2 ... 8 tiles per CTA at these sizes. The list must be the
    ascending list of the points whose error (from the independent per-point error kernel) is below the threshold,
    the score the sum of min(e, thr), twice in a row (epoch-tagged status words) and for a threshold that keeps
    nothing / everything."""
    rng = np.random.default_rng(n)
    inst = synth.pose_instance(55, 4000, 4000)
    f = inst["focal"]
    reps = n // 8000 + 1
    p2 = np.tile(inst["points2D"], (reps, 1))[:n] + rng.normal(0, 0.5, (n, 2))
    p3 = np.tile(inst["points3D"], (reps, 1))[:n]
    sc = ctx.pose_scorer(p2, p3, f, f)
    m = synth.pose_to_model(inst["R_gt"], inst["c_gt"])
    err = sc.eval_errors(m)
    for thr in (144.0, 1e-12, 1e12):
        want = np.nonzero(err < thr)[0]
        for _ in range(2):
            lists, counts, score = sc.get_inliers(m, thr, thr)
            assert counts[0] == len(want)
            assert np.array_equal(lists[0], want.astype(np.int32))
            np.testing.assert_allclose(score, np.minimum(err, thr).sum(), rtol=1e-12)
    sc.close()


def test_inlier_pass_one_wave_two_data_types(ctx):
    """The same for the hybrid line scorer (two data types, per-type thresholds and weights, the type boundary inside
    a CTA's tile range)."""
    n0, n1 = 700_000, 650_000
    xy0, p1 = synth.hybrid_line_instance(9, n0 // 2, n0 - n0 // 2, n1 // 2, n1 - n1 // 2)
    sc = ctx.hybrid_line_scorer(xy0, p1)
    m = np.array([0.6, -0.8, 0.1])
    e0, e1 = sc.eval_errors(m, type_=0), sc.eval_errors(m, type_=1)
    thr, w = [0.5, 0.02], [2.0, 0.5]
    lists, counts, score = sc.get_inliers(m, thr, thr, weights=w)
    assert np.array_equal(lists[0], np.nonzero(e0 < thr[0])[0].astype(np.int32))
    assert np.array_equal(lists[1], np.nonzero(e1 < thr[1])[0].astype(np.int32))
    want = w[0] * np.minimum(e0, thr[0]).sum() + w[1] * np.minimum(e1, thr[1]).sum()
    np.testing.assert_allclose(score, want, rtol=1e-12)
    sc.close()
