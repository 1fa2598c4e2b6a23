"""GPU tests of the in-library communicator (rlb_comm_*, SURVEY.md 8(e) hypothesis sharding):
one-rank communicators must equal the plain scorer bit for bit; the cross-rank tie-break
(lowest global id among equal scores, RansacLib/ransac.h:195,288) is checked with the ranks'
shards emulated one after the other on this GPU; with >= 2 GPUs visible the real NCCL path runs."""
import numpy as np
import pytest

import oracle_lib as ol
from ransaclib_b200 import sharding, synth
from ransaclib_b200.sampling import UniformSampling

pytestmark = pytest.mark.gpu
THR = 144.0


def rlb():
    import ransaclib_b200 as m

    return m


@pytest.fixture(scope="module")
def case():
    inst = synth.pose_instance(31, 1500, 1300)
    samples = UniformSampling(9, 2800, 4).sample_fast(2048)
    return inst, inst["focal"], samples


@pytest.mark.parametrize("form", ["from_rank", "local"])
@pytest.mark.parametrize("precision", ["f64", "f32"])
def test_one_rank_communicator_equals_the_scorer(ctx, case, form, precision):
    inst, f, samples = case
    m = rlb()
    prec = m.RLB_F64 if precision == "f64" else m.RLB_F32
    sc = ctx.pose_scorer(inst["points2D"], inst["points3D"], f, f)
    ref = sc.hypothesize_and_score(samples, THR, precision=prec)
    comm = m.Comm.from_rank(ctx, 1, 0, None) if form == "from_rank" else m.Comm.local(1)
    assert comm.nranks == 1 and comm.num_local == 1 and comm.rank(0) == 0
    cs = comm.pose_scorer(inst["points2D"], inst["points3D"], f, f, root=0 if form == "from_rank" else -1)
    out = cs.hypothesize_and_score(samples, THR, precision=prec, want_scores=True)
    # the communicator's scorer is built from the DEVICE copy of the points (rlb_scorer_create_pose_dev):
    # same centre, same repack, so even the f32 scores are the same bits
    assert out["best_id"] == ref["best_id"] and out["best_score"] == ref["best_score"]
    assert out["best_model"].tolist() == ref["best_model"].tolist()
    assert out["scores"].tolist() == ref["scores"].tolist()
    assert out["counts"].tolist() == ref["counts"].tolist()
    inl, score = cs.get_inliers(out["best_model"], THR, THR)
    want = ol.pose_inliers(ol.oracle(), f, f, inst["points2D"], inst["points3D"], out["best_model"], THR)
    assert inl.tolist() == want.tolist()
    assert comm.collective_count == 0  # one rank: nothing to gather
    cs.close()
    comm.close()
    sc.close()


def test_cross_rank_tie_break_with_emulated_ranks(ctx, case):
    """The best hypothesis appears TWICE (ids a < b) in different shards with bit-identical scores, and a
    third copy sits in the last shard: the winner record must carry the LOWEST global id for every world
    size, exactly as the sequential loop keeps the first of equal scores (strict <, ransac.h:195)."""
    import torch

    inst, f, samples = case
    sc = ctx.pose_scorer(inst["points2D"], inst["points3D"], f, f)
    base = sc.hypothesize_and_score(samples, THR)
    best = base["best_id"]
    s = samples.copy()
    k_total = len(s)
    a, b, c = 700, 1500, k_total - 1
    assert best not in (a, b, c)
    s[a] = s[b] = s[c] = samples[best]
    s[best] = samples[(best + 1) % k_total]  # the original goes away; a is now the first of three equal minima
    ref = sc.hypothesize_and_score(s, THR)
    assert ref["best_id"] == a and ref["scores"][a] == ref["scores"][b] == ref["scores"][c]
    dev = torch.device("cuda", 0)
    for world in (2, 3, 5, 8):
        recs = torch.zeros((world, 14), dtype=torch.float64, device=dev)
        owners = set()
        for rank in range(world):
            lo, hi = sharding.hypothesis_range(k_total, rank, world)
            for i in (a, b, c):
                if lo <= i < hi:
                    owners.add(rank)
            k = hi - lo
            d_s = torch.from_numpy(np.ascontiguousarray(s[lo:hi])).to(dev)
            d_m = torch.empty((k, 12), dtype=torch.float64, device=dev)
            d_v = torch.empty(k, dtype=torch.uint8, device=dev)
            torch.cuda.synchronize()
            sc.sharded_round_dev(d_s.data_ptr(), k, THR, None, lo, d_m.data_ptr(), d_v.data_ptr(), 0, 0,
                                 recs[rank].data_ptr())
            ctx.synchronize()
        assert len(owners) >= 2, "the equal minima must sit on different ranks"
        r = recs.cpu().numpy()
        assert (r[:, 0] == ref["best_score"]).sum() == len(owners)  # every owner reports the same score
        # the ranks' records in any arrival order (an all_gather is rank-ordered; the rule must not rely on it)
        for perm in (np.arange(world), np.arange(world)[::-1].copy()):
            g = recs[torch.from_numpy(perm).to(dev)].contiguous()
            win = torch.zeros(14, dtype=torch.float64, device=dev)
            ctx.select_winner_dev(g.data_ptr(), world, 14, win.data_ptr())
            ctx.synchronize()
            w = win.cpu().numpy()
            assert int(w[1]) == a and w[0] == ref["best_score"], (world, perm.tolist(), w[:2])
            assert w[2:].tolist() == ref["best_model"].tolist()
    sc.close()


def test_no_model_anywhere_and_more_ranks_than_hypotheses(ctx):
    """A round in which no sample yields a model reports id -1; a 1-rank communicator with k_total = 1 works."""
    m = rlb()
    inst = synth.pose_instance(5, 0, 60)  # outliers only: P3P + 4th-point test rejects (almost) everything
    f = inst["focal"]
    samples = UniformSampling(3, 60, 4).sample_fast(64)
    sc = ctx.pose_scorer(inst["points2D"], inst["points3D"], f, f)
    ref = sc.hypothesize_and_score(samples, 1e-6)  # tiny threshold: nothing passes the 4th-point test
    comm = m.Comm.from_rank(ctx, 1, 0, None)
    cs = comm.pose_scorer(inst["points2D"], inst["points3D"], f, f)
    out = cs.hypothesize_and_score(samples, 1e-6)
    assert out["best_id"] == ref["best_id"]
    one = cs.hypothesize_and_score(samples[:1], THR)
    assert one["best_id"] in (-1, 0)
    cs.close()
    comm.close()
    sc.close()


def test_two_gpus_one_process(case):
    """rlb_comm_create_local over 2 GPUs: ncclBroadcast of the points, one ncclAllGather per round, winner ==
    the single-GPU winner. Skipped on a one-GPU box (the driver's scaling run covers N = 2, 4, 8)."""
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    m = rlb()
    inst, f, samples = case
    c0 = m.Context(0)
    sc = c0.pose_scorer(inst["points2D"], inst["points3D"], f, f)
    ref = sc.hypothesize_and_score(samples, THR)
    comm = m.Comm.local(2)
    try:
        # (a shard of 1024 hypotheses and the whole batch of 2048 are planned with different point-axis splits:
        # the same pairs summed in another grouping, hence 1e-12 and not bit equality for the scores)
        def same(out):
            assert out["best_id"] == ref["best_id"]
            assert abs(out["best_score"] - ref["best_score"]) <= 1e-12 * abs(ref["best_score"])
            big = ref["scores"] > 1e300
            assert ((out["scores"] > 1e300) == big).all()
            np.testing.assert_allclose(out["scores"][~big], ref["scores"][~big], rtol=1e-12)

        cs = comm.pose_scorer(inst["points2D"], inst["points3D"], f, f)
        same(cs.hypothesize_and_score(samples, THR, want_scores=True))
        assert comm.collective_count == 3  # the two point arrays broadcast, one all_gather
        cs.close()
        # every process holds the arrays: each GPU uploads half of them, the halves cross NVLink - equal halves
        # (n = 2800) by two in-place ncclAllGathers
        cs = comm.pose_scorer(inst["points2D"], inst["points3D"], f, f, root=-1)
        same(cs.hypothesize_and_score(samples, THR, want_scores=True))
        assert comm.collective_count == 3 + 2 + 1
        cs.close()
        # ... uneven ones (n = 2799) by one broadcast per array and rank; the samples never name the last point
        before = comm.collective_count
        odd = np.minimum(samples, 2798)
        ref_odd = c0.pose_scorer(inst["points2D"][:-1], inst["points3D"][:-1], f, f)
        want = ref_odd.hypothesize_and_score(odd, THR)
        cs = comm.pose_scorer(inst["points2D"][:-1], inst["points3D"][:-1], f, f, root=-1)
        got = cs.hypothesize_and_score(odd, THR, want_scores=True)
        assert comm.collective_count == before + 2 * 2 + 1
        assert got["best_id"] == want["best_id"]
        assert abs(got["best_score"] - want["best_score"]) <= 1e-12 * abs(want["best_score"])
        cs.close()
        ref_odd.close()
    finally:
        comm.close()
        sc.close()
        c0.close()


@pytest.mark.parametrize("gpus", [1, 2])
def test_drop_in_layer_shards_large_batches(gpus, monkeypatch):
    """LocallyOptimizedMSAC over CalibratedAbsolutePoseEstimator with RLB_NUM_GPUS=G: the data set is replicated by
    the library's communicator (gpu::DeviceScorer::InitPoseMulti), batches of >= 2048 hypotheses are sharded and
    settled by one ncclAllGather each - and the run is the single-GPU run, statistic for statistic."""
    import torch

    from ransaclib_b200 import host as H

    if torch.cuda.device_count() < gpus:
        pytest.skip("needs %d GPUs" % gpus)
    inst = synth.pose_instance(4242, 5000, 5000)
    f = inst["focal"]
    opts = H.LORansacOptions(squared_inlier_threshold_=144.0, random_seed_=42, min_num_iterations_=10000,
                             max_num_iterations_=10000)
    monkeypatch.delenv("RLB_NUM_GPUS", raising=False)
    m1, st1 = H.estimate_pose(opts, f, f, inst["points2D"], inst["points3D"])
    monkeypatch.setenv("RLB_NUM_GPUS", str(gpus))
    mg, stg = H.estimate_pose(opts, f, f, inst["points2D"], inst["points3D"])
    assert stg["num_iterations"] == st1["num_iterations"] == 10000
    assert mg.tolist() == m1.tolist()
    for key in ("best_num_inliers", "number_lo_iterations", "best_model_score"):
        assert stg[key] == st1[key], key
    assert stg["inlier_indices"].tolist() == st1["inlier_indices"].tolist()
