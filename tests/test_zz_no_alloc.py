"""rlb_get_inliers_noalloc_dev (the allocation-free variant of the K = 1 pass; no counterpart in the reference)
against rlb_get_inliers_dev through the C ABI: same header (score,
per-type counts) and the same ascending inlier lists, on two calls in a row with different models, so that whatever
the shared static scratch keeps from one call is seen by the next. Counts and lists must be identical; the score is
compared with the 1e-9 relative tolerance of the f64 path (it is in fact the same bits: same kernel, same order).

This file sorts last on purpose: a failure here hides no other test."""
import numpy as np
import pytest

from ransaclib_b200 import _capi, synth
from ransaclib_b200.sampling import UniformSampling

pytestmark = pytest.mark.gpu
F64_RTOL = 1e-9


def _run(sc, fn, d_model, thr_inl, thr_score, weights, n_total):
    """One pass through `fn` (a bound Scorer method) -> (score, counts[2], idx list)."""
    import torch

    dev = d_model.device
    d_idx = torch.full((max(n_total, 1),), -7, dtype=torch.int32, device=dev)
    d_hdr = torch.zeros(4, dtype=torch.float64, device=dev)  # {double score; int32 count[2]; int32 pad[2]}
    fn(d_model.data_ptr(), thr_inl, thr_score, weights, d_idx.data_ptr(), d_hdr.data_ptr())
    sc.ctx.synchronize()
    hdr = d_hdr.cpu().numpy()
    counts = hdr.view(np.int32)[2:4].copy()
    return float(hdr[0]), counts, d_idx.cpu().numpy()


def _compare(sc, d_model, thr_inl, thr_score, weights, n_total):
    want = _run(sc, sc.get_inliers_dev, d_model, thr_inl, thr_score, weights, n_total)
    got = _run(sc, sc.get_inliers_noalloc_dev, d_model, thr_inl, thr_score, weights, n_total)
    assert got[1].tolist() == want[1].tolist()
    total = int(want[1][1])  # count[1] = inliers over all types
    assert total > 0
    assert got[2][:total].tolist() == want[2][:total].tolist()
    np.testing.assert_allclose(got[0], want[0], rtol=F64_RTOL)
    return total


def test_noalloc_inlier_pass_equals_the_existing_entry_point_twice_in_a_row(ctx):
    import torch

    dev = torch.device("cuda", 0)
    # pose, N = 60,000 (59 tiles of 1024 points): two different models, one after the other, then the first again
    inst = synth.pose_instance(321, 36000, 24000)
    f = inst["focal"]
    sc = ctx.pose_scorer(inst["points2D"], inst["points3D"], f, f)
    samples = UniformSampling(4, 60000, 4).sample_fast(256)
    r = sc.hypothesize_and_score(samples, 144.0)
    ok = np.flatnonzero(r["valid"] == 1)
    order = ok[np.argsort(r["scores"][ok])]
    models = [r["models"][order[0]], r["models"][order[len(order) // 2]], r["models"][order[0]]]
    totals = []
    for m in models:
        d_model = torch.from_numpy(np.ascontiguousarray(m)).to(dev)
        totals.append(_compare(sc, d_model, 144.0, 144.0, None, 60000))
    assert totals[0] == totals[2] and totals[0] != totals[1]
    # inlier threshold != score threshold (the LO pattern), and no score at all
    d_model = torch.from_numpy(np.ascontiguousarray(models[0])).to(dev)
    _compare(sc, d_model, 288.0, 144.0, None, 60000)
    want = _run(sc, sc.get_inliers_dev, d_model, 144.0, None, None, 60000)
    got = _run(sc, sc.get_inliers_noalloc_dev, d_model, 144.0, None, None, 60000)
    assert got[1].tolist() == want[1].tolist() and got[2][:want[1][1]].tolist() == want[2][:want[1][1]].tolist()
    sc.close()

    # hybrid line: two data types, per-type thresholds and weights, a different tile count right after
    xy0, p1 = synth.hybrid_line_instance(17, 4000, 3000, 2500, 2600)
    hs = ctx.hybrid_line_scorer(xy0, p1)
    s2 = UniformSampling(5, len(xy0), 2).sample_fast(64)
    hr = hs.hypothesize_and_score(s2, [1e-4, 1e-4], weights=[2.0, 0.5])
    for m in (hr["best_model"], hr["models"][int(np.flatnonzero(hr["valid"] == 1)[-1])]):
        d_model = torch.from_numpy(np.ascontiguousarray(m)).to(dev)
        want = _run(hs, hs.get_inliers_dev, d_model, [1e-4, 2e-4], [1e-4, 1e-4], [2.0, 0.5], len(xy0) + len(p1))
        got = _run(hs, hs.get_inliers_noalloc_dev, d_model, [1e-4, 2e-4], [1e-4, 1e-4], [2.0, 0.5],
                   len(xy0) + len(p1))
        assert got[1].tolist() == want[1].tolist()
        assert got[2][:want[1][1]].tolist() == want[2][:want[1][1]].tolist()
        np.testing.assert_allclose(got[0], want[0], rtol=F64_RTOL)
    hs.close()


def test_noalloc_rejects_sizes_beyond_its_fixed_scratch(ctx):
    import torch

    from ransaclib_b200 import RlbError

    n = _capi.NOALLOC_MAX_POINTS + 1
    rng = np.random.default_rng(3)
    xy = rng.uniform(-1.0, 1.0, (n, 2))
    sc = ctx.line_scorer(xy)
    dev = torch.device("cuda", 0)
    d_model = torch.tensor([0.6, 0.8, 0.1], dtype=torch.float64, device=dev)
    d_hdr = torch.zeros(4, dtype=torch.float64, device=dev)
    with pytest.raises(RlbError, match="RLB_NOALLOC_MAX_POINTS"):
        sc.get_inliers_noalloc_dev(d_model.data_ptr(), 1e-4, 1e-4, None, 0, d_hdr.data_ptr())
    # the existing entry point still serves that size
    sc.get_inliers_dev(d_model.data_ptr(), 1e-4, 1e-4, None, 0, d_hdr.data_ptr())
    ctx.synchronize()
    assert d_hdr.cpu().numpy().view(np.int32)[3] >= 0
    sc.close()
    # exactly the maximum is accepted
    sc = ctx.line_scorer(xy[:_capi.NOALLOC_MAX_POINTS])
    d_idx = torch.zeros(_capi.NOALLOC_MAX_POINTS, dtype=torch.int32, device=dev)
    a = _run(sc, sc.get_inliers_dev, d_model, 1e-4, 1e-4, None, _capi.NOALLOC_MAX_POINTS)
    b = _run(sc, sc.get_inliers_noalloc_dev, d_model, 1e-4, 1e-4, None, _capi.NOALLOC_MAX_POINTS)
    assert a[1].tolist() == b[1].tolist() and a[2][:a[1][1]].tolist() == b[2][:a[1][1]].tolist()
    del d_idx
    sc.close()
