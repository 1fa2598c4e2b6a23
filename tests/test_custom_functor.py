"""User-defined EvaluateModelOnPoint functors (rlb_scorer_create_custom): the NVRTC compile step runs
without a device, the scoring tests need one."""
import numpy as np
import pytest

LINE_SRC = """
__device__ double rlb_eval(const double* m, const double* p) {
  const double r = m[0] * p[0] + m[1] * p[1] + m[2];   // line_estimator.cc:110-114
  return r * r;
}
"""

CIRCLE_SRC = """
__device__ double rlb_eval(const double* m, const double* p) {
  const double dx = p[0] - m[0], dy = p[1] - m[1];
  const double r = sqrt(dx * dx + dy * dy) - m[2];
  return r * r;
}
"""

PLANE_SRC = """
__device__ double rlb_eval(const double* m, const double* p) {
  const double r = m[0] * p[0] + m[1] * p[1] + m[2] * p[2] + m[3];
  return r * r;
}
"""


def test_compile_check_without_a_device():
    import ransaclib_b200 as rlb

    assert rlb.custom_compile_check(LINE_SRC, 2, 3) > 1000
    assert rlb.custom_compile_check(PLANE_SRC, 3, 4) > 1000
    with pytest.raises(Exception) as ei:
        rlb.custom_compile_check("__device__ double rlb_eval(const double* m, const double* p) { return nope; }", 2, 3)
    assert "nope" in str(ei.value)  # the compiler log comes back
    for dims, md in ((0, 3), (9, 3), (2, 0), (2, 13)):
        with pytest.raises(Exception):
            rlb.custom_compile_check(LINE_SRC, dims, md)


def _line_data(seed, n):
    rng = np.random.default_rng(seed)
    x = rng.uniform(0, 1, n)
    y = 0.3 * x + 0.1 + rng.normal(0, 0.003, n)
    out = rng.uniform(0, 1, (n // 3, 2))
    return np.concatenate([np.stack([x, y], 1), out])


def _line_models(seed, k):
    rng = np.random.default_rng(seed)
    th = rng.uniform(0, np.pi, k)
    a, b = np.cos(th), np.sin(th)
    a[0], b[0] = 0.3 / np.hypot(0.3, 1), -1 / np.hypot(0.3, 1)
    c = rng.uniform(-0.5, 0.5, k)
    c[0] = 0.1 / np.hypot(0.3, 1)
    return np.stack([a, b, c], 1)


@pytest.mark.gpu
@pytest.mark.parametrize("n", [7, 128, 1000, 5003])
def test_custom_line_equals_builtin_line(ctx, n):
    xy = _line_data(1, n)
    models = _line_models(2, 700)
    valid = np.ones(len(models), np.uint8)
    valid[5] = 0
    models[9, 1] = np.nan
    ref = ctx.line_scorer(xy)
    cus = ctx.custom_scorer(xy, 3, LINE_SRC)
    assert cus.kind == 3 and cus.model_size == 3 and cus.num_data == [len(xy)]
    thr = 1e-4
    s0, c0, b0, bs0 = ref.score_models(models, thr, valid=valid)
    s1, c1, b1, bs1 = cus.score_models(models, thr, valid=valid)
    ok = np.isfinite(s0) & (valid == 1)
    np.testing.assert_allclose(s1[ok], s0[ok], rtol=1e-9)
    assert (c1[ok] == c0[ok]).all() and b1 == b0
    assert s1[5] == np.finfo(np.float64).max and c1[5] == 0 and np.isnan(s1[9])
    s2, c2, b2, _ = cus.score_models(models, thr, valid=valid, want_counts=False)
    assert c2 is None and b2 == b0 and s2[ok].tolist() == s1[ok].tolist()
    l0, n0, sc0 = ref.get_inliers(models[b0], thr, thr)
    l1, n1, sc1 = cus.get_inliers(models[b0], thr, thr)
    assert l1[0].tolist() == l0[0].tolist() and n1.tolist() == n0.tolist()
    np.testing.assert_allclose(sc1, sc0, rtol=1e-9)
    cus.close()
    ref.close()


@pytest.mark.gpu
def test_custom_circle_and_plane_against_numpy(ctx):
    rng = np.random.default_rng(3)
    n = 3001
    ang = rng.uniform(0, 2 * np.pi, n)
    pts = np.stack([1.5 + 2.0 * np.cos(ang), -0.5 + 2.0 * np.sin(ang)], 1) + rng.normal(0, 0.01, (n, 2))
    pts[::4] = rng.uniform(-3, 4, (len(pts[::4]), 2))
    models = np.stack([1.5 + rng.normal(0, 0.2, 513), -0.5 + rng.normal(0, 0.2, 513), 2.0 + rng.normal(0, 0.1, 513)], 1)
    models[17] = [1.5, -0.5, 2.0]
    thr = 0.03 ** 2
    sc = ctx.custom_scorer(pts, 3, CIRCLE_SRC)
    dx, dy = pts[None, :, 0] - models[:, None, 0], pts[None, :, 1] - models[:, None, 1]
    d = np.sqrt(dx * dx + dy * dy) - models[:, None, 2]  # the functor's own sequence of operations
    e = d * d
    want_s = np.minimum(e, thr).sum(1)
    want_c = (e < thr).sum(1)
    s, c, bid, bsc = sc.score_models(models, thr)
    np.testing.assert_allclose(s, want_s, rtol=1e-9)
    near = (np.abs(e - thr) <= 1e-9 * thr).sum(1)
    assert (np.abs(c - want_c) <= near).all()
    assert bid == int(np.argmin(want_s)) and bsc == s[bid]
    lists, counts, score = sc.get_inliers(models[bid], thr, thr)
    want_idx = np.nonzero(e[bid] < thr)[0]
    assert lists[0].tolist() == want_idx.tolist() and counts[0] == len(want_idx)
    np.testing.assert_allclose(score, want_s[bid], rtol=1e-9)
    np.testing.assert_allclose(sc.eval_errors(models[bid]), e[bid], rtol=1e-9, atol=1e-300)
    sc.close()

    p3 = rng.normal(0, 1, (777, 3))
    p3[:, 2] = 0.2 * p3[:, 0] - 0.1 * p3[:, 1] + 0.5 + rng.normal(0, 0.005, 777)
    planes = rng.normal(0, 1, (600, 4))
    planes[3] = np.array([0.2, -0.1, -1.0, 0.5]) / np.linalg.norm([0.2, -0.1, -1.0])
    ps = ctx.custom_scorer(p3, 4, PLANE_SRC)
    r = p3 @ planes[:, :3].T + planes[:, 3]
    e = (r * r).T
    s, c, bid, _ = ps.score_models(planes, 1e-4)
    np.testing.assert_allclose(s, np.minimum(e, 1e-4).sum(1), rtol=1e-9)
    assert bid == 3
    ps.close()


POSE_SRC = """
__device__ double rlb_eval(const double* m, const double* p) {   // p = (u, v, X, Y, Z), focal 185
  const double d0 = p[2] - m[9], d1 = p[3] - m[10], d2 = p[4] - m[11];
  const double pz = m[2] * d0 + m[5] * d1 + m[8] * d2;
  if (pz < 0.0) return 1.7976931348623157e308;
  const double px = m[0] * d0 + m[3] * d1 + m[6] * d2;
  const double py = m[1] * d0 + m[4] * d1 + m[7] * d2;
  const double du = DIV(px, pz) * 185.0 - p[0], dv = DIV(py, pz) * 185.0 - p[1];
  return du * du + dv * dv;
}
"""


@pytest.mark.gpu
def test_rlb_rcp_helper(ctx):
    """rlb_rcp(x), the branch-free f64 reciprocal the prelude offers to functor authors (a / b in f64 compiles to a
    call with a slow path): within an ulp of IEEE division, so the sweep with it agrees with the sweep that divides
    to ~1e-15 and selects the same model; and the plain form is the reference's arithmetic (oracle, 1e-9)."""
    import oracle_lib as ol
    from ransaclib_b200 import synth

    inst = synth.pose_instance(5, 2000, 2000, focal=185.0)
    rows = np.ascontiguousarray(np.concatenate([inst["points2D"], inst["points3D"]], axis=1))
    models = synth.perturbed_poses(3, inst["R_gt"], inst["c_gt"], 700)
    a = ctx.custom_scorer(rows, 12, "#define DIV(x, y) ((x) / (y))\n" + POSE_SRC)
    b = ctx.custom_scorer(rows, 12, "#define DIV(x, y) ((x) * rlb_rcp(y))\n" + POSE_SRC)
    sa, ca, ba, _ = a.score_models(models, 144.0)
    sb, cb, bb, _ = b.score_models(models, 144.0)
    ws, wc = ol.pose_score(ol.oracle(), 185.0, 185.0, inst["points2D"], inst["points3D"], models, 144.0)
    np.testing.assert_allclose(sa, ws, rtol=1e-9)
    assert (ca == wc).all()
    np.testing.assert_allclose(sb, sa, rtol=1e-12)
    assert np.abs(cb - ca).max() <= 1 and bb == ba
    a.close()
    b.close()


@pytest.mark.gpu
def test_custom_rejects_device_solvers_and_bad_source(ctx):
    xy = _line_data(4, 100)
    cus = ctx.custom_scorer(xy, 3, LINE_SRC)
    with pytest.raises(Exception):
        cus.solve_line2pt(np.array([[0, 1]], np.int32))
    cus.close()
    with pytest.raises(Exception) as ei:
        ctx.custom_scorer(xy, 3, "__device__ double rlb_eval(const double* m, const double* p) { return oops; }")
    assert "oops" in str(ei.value)


# ------------------------------------------------------------------ two data types (the HybridSolver concept)
HYBRID_LINE_SRC = """
__device__ double rlb_eval_t(const double* m, int type, const double* p) {
  const double r = m[0] * p[0] + m[1] * p[1] + m[2] * 1.0;   // hybrid_line_estimator.cc:100-109
  return r * r;
}
"""

# a functor that really depends on the data type: type 1 also penalises the angle between the line's normal and
# the point's own normal
HYBRID_NORMAL_SRC = """
__device__ double rlb_eval_t(const double* m, int type, const double* p) {
  const double r = m[0] * p[0] + m[1] * p[1] + m[2];
  if (type == 0) return r * r;
  const double c = m[0] * p[2] + m[1] * p[3];
  return r * r + 0.01 * (1.0 - c * c);
}
"""


def test_hybrid_compile_check_without_a_device():
    import ransaclib_b200 as rlb

    assert rlb.custom_compile_check(HYBRID_LINE_SRC, 4, 3) > 1000


@pytest.mark.gpu
@pytest.mark.parametrize("n0,n1", [(3000, 2500), (1024, 1), (5, 2049), (0, 700)])
def test_hybrid_custom_line_equals_builtin_hybrid_line(ctx, n0, n1):
    """rlb_scorer_create_custom_hybrid with the hybrid line residual against the built-in hybrid line scorer:
    weighted scores, per-type counts, per-type ascending inlier lists, errors of both types."""
    from ransaclib_b200 import synth

    xy0, p1 = synth.hybrid_line_instance(77, max(n0 - n0 // 3, 0), n0 // 3, max(n1 - n1 // 3, 0), n1 // 3)
    xy0, p1 = xy0[:n0], p1[:n1]
    assert len(xy0) == n0 and len(p1) == n1
    models = _line_models(5, 300)
    thr, w = [1e-4, 4e-4], [2.0, 0.5]
    sb = ctx.hybrid_line_scorer(xy0, p1)
    sc = ctx.custom_hybrid_scorer(xy0, p1, 3, HYBRID_LINE_SRC)
    assert sc.num_types == 2 and sc.num_data == [n0, n1]
    s0, c0, b0, _ = sb.score_models(models, thr, weights=w)
    s1, c1, b1, _ = sc.score_models(models, thr, weights=w)
    np.testing.assert_allclose(s1, s0, rtol=1e-9)
    assert c1.tolist() == c0.tolist() and b1 == b0
    for m in (models[b0], models[1]):
        l0, k0, sc0 = sb.get_inliers(m, thr, thr, weights=w)
        l1, k1, sc1 = sc.get_inliers(m, thr, thr, weights=w)
        assert k1.tolist() == k0.tolist()
        assert [a.tolist() for a in l1] == [a.tolist() for a in l0]
        np.testing.assert_allclose(sc1, sc0, rtol=1e-9)
        for t, n in ((0, n0), (1, n1)):
            if n:
                np.testing.assert_allclose(sc.eval_errors(m, type_=t), sb.eval_errors(m, type_=t), rtol=1e-9, atol=1e-18)


@pytest.mark.gpu
def test_hybrid_custom_functor_sees_the_data_type(ctx):
    rng = np.random.default_rng(3)
    d0 = rng.uniform(-1, 1, (1500, 2))
    d1 = np.concatenate([rng.uniform(-1, 1, (900, 2)), rng.normal(size=(900, 2))], axis=1)
    d1[:, 2:] /= np.linalg.norm(d1[:, 2:], axis=1, keepdims=True)
    models = _line_models(8, 64)
    thr, w = [0.01, 0.02], [1.0, 3.0]
    sc = ctx.custom_hybrid_scorer(d0, d1, 3, HYBRID_NORMAL_SRC)
    s, c, bid, _ = sc.score_models(models, thr, weights=w)
    for k in (0, 7, bid):
        m = models[k]
        e0 = (d0 @ m[:2] + m[2]) ** 2
        r1 = d1[:, :2] @ m[:2] + m[2]
        e1 = r1 ** 2 + 0.01 * (1.0 - (d1[:, 2:] @ m[:2]) ** 2)
        want = w[0] * np.minimum(e0, thr[0]).sum() + w[1] * np.minimum(e1, thr[1]).sum()
        np.testing.assert_allclose(s[k], want, rtol=1e-9)
        assert c[k] == int((e0 < thr[0]).sum() + (e1 < thr[1]).sum())
        lists, counts, _ = sc.get_inliers(m, thr)
        assert lists[0].tolist() == np.nonzero(e0 < thr[0])[0].tolist()
        assert lists[1].tolist() == np.nonzero(e1 < thr[1])[0].tolist()


@pytest.mark.gpu
def test_hybrid_solver_with_a_custom_functor_runs_like_the_builtin_one():
    """HybridLocallyOptimizedMSAC over examples/hybrid_custom_line_estimator.h (EvaluateModelOnPoint handed over as
    source) against the same run over the built-in HybridLineEstimator: same iterations, solver choices, LO count,
    per-type inlier lists and model."""
    from ransaclib_b200 import host as H
    from ransaclib_b200 import synth

    for n, seed in [(100, 42), (20000, 6)]:
        n_out = n // 2
        xy0, p1 = synth.hybrid_line_instance(300 + seed, n - n_out, n_out, n - n_out, n_out)
        opts = H.HybridLORansacOptions(squared_inlier_thresholds_=[1e-4, 1e-4], data_type_weights_=[2.0, 0.5],
                                       random_seed_=seed, min_num_iterations_=100, max_num_iterations_=10000,
                                       max_num_iterations_per_solver_=1000)
        m0, st0 = H.estimate_hybrid_line(opts, xy0, p1, [0.2, 0.8])
        m1, st1 = H.estimate_hybrid_line(opts, xy0, p1, [0.2, 0.8], custom_functor=True)
        for key in ("num_iterations_total", "num_iterations_per_solver", "best_num_inliers", "best_solver_type",
                    "number_lo_iterations", "inlier_ratios"):
            assert st1[key] == st0[key], key
        np.testing.assert_allclose(st1["best_model_score"], st0["best_model_score"], rtol=1e-9)
        np.testing.assert_allclose(m1, m0, rtol=1e-9, atol=1e-12)
        for t in (0, 1):
            assert st1["inlier_indices"][t].tolist() == st0["inlier_indices"][t].tolist()
