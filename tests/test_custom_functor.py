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
