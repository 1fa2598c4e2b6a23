"""CPU tests: the restated oracle (oracle/liboracle.so, built from committed sources)
against the golden vectors produced by the reference's own headers (oracle/_ref) and
against OpenCV's P3P. These pin the oracle; the GPU parity tests then compare the
CUDA path with the pinned oracle."""
import ctypes as C
import json
import os

import numpy as np
import pytest

import oracle_lib as ol

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def G():
    with open(os.path.join(HERE, "golden", "reference_vectors.json")) as fh:
        return json.load(fh)


@pytest.fixture(scope="module")
def O():
    return ol.oracle()


def test_survey_known_answers(O):
    # vectors produced by the unmodified reference headers.
    assert ol.sample_uniform(O, 0, 10000, 4, 3).tolist() == [[5488, 5928, 7151, 8442], [6027, 8579, 5448, 8472],
                                                           [4236, 6235, 6458, 3843]]
    assert ol.sample_uniform(O, 42, 10000, 4, 3).tolist() == [[3745, 7965, 9507, 1834], [7319, 7796, 5986, 5968],
                                                            [1560, 4458, 1559, 999]]
    assert ol.sample_uniform(O, 42, 5, 4, 3).tolist() == [[1, 4, 0, 3], [3, 4, 0, 1], [0, 2, 1, 3]]
    assert O.num_required_iterations(0.5, 1e-4, 4, 100, 10000) == 144
    assert O.num_required_iterations(0.1, 1e-4, 4, 100, 1000000) == 92100


def test_uniform_sampling(G, O):
    for c in G["uniform_sampling"]:
        got = ol.sample_uniform(O, c["seed"], c["n"], c["s"], c["k"])
        assert got.tolist() == c["samples"]


def test_hybrid_sampling(G, O):
    for c in G["hybrid_sampling"]:
        k = len(c["sizes"])
        o = np.empty((k, 2), np.int32)
        O.sample_hybrid(c["seed"], 2, np.asarray(c["num_data"], np.int32), k,
                        np.asarray(c["sizes"], np.int32).reshape(-1), 2, o.reshape(-1))
        assert o.tolist() == c["samples"]


def test_hybrid_biased_sampling(G, O):
    """HybridBiasedSampling (reference hybrid_sampling.h:146-223): zero weights never drawn, discrete_distribution
    stream of the reference, the all-candidates short-cut."""
    for c in G["hybrid_biased_sampling"]:
        k = len(c["sizes"])
        stride = len(c["samples"][0])
        o = np.empty((k, stride), np.int32)
        w = np.asarray(c["weights"][0] + c["weights"][1], np.float64)
        O.sample_hybrid_biased(c["seed"], 2, np.asarray(c["num_data"], np.int32), w, k,
                               np.asarray(c["sizes"], np.int32).reshape(-1), stride, o.reshape(-1))
        assert o.tolist() == c["samples"]
        w0, w1 = c["weights"]
        for row, sz in zip(c["samples"], c["sizes"]):
            assert all(w0[i] > 0 for i in row[:sz[0]]) and all(w1[i] > 0 for i in row[sz[0]:sz[0] + sz[1]])


def test_template_layer_samplers_match_the_reference(G):
    """The drop-in headers' own samplers (include/RansacLib/sampling.h, hybrid_sampling.h), called through the host
    library without a device, reproduce the reference's sample streams."""
    from ransaclib_b200 import host as H

    for c in G["uniform_sampling"]:
        assert H.sample_uniform(c["seed"], c["n"], c["s"], c["k"]).tolist() == c["samples"]
    for key in ("hybrid_sampling", "hybrid_biased_sampling"):
        for c in G[key]:
            got = H.sample_hybrid(c["seed"], c["num_data"], c["sizes"], c.get("weights"))
            flat = [[i for t in draw for i in t] for draw in got]
            want = [[i for i in row if i >= 0] for row in c["samples"]]
            assert flat == want, key


def test_num_required_iterations(G, O):
    for c in G["num_required_iterations"]:
        assert O.num_required_iterations(*c["args"]) == c["value"]
    for c in G["num_required_iterations_hybrid"]:
        v = O.num_required_iterations_hybrid(2, np.asarray(c["ratios"], np.float64), 1e-4,
                                             np.asarray(c["sizes"], np.int32), c["min"], c["max"])
        assert v == c["value"]


def test_shuffle_consumes_same_draws(G, O):
    for c in G["shuffle_and_resize"]:
        o = np.empty(c["target"], np.int32)
        nxt = C.c_uint32()
        O.shuffle_and_resize(c["seed"], c["n"], c["target"], o, C.byref(nxt))
        assert o.tolist() == c["out"]
        assert nxt.value == c["next_draw"]


def test_drivers_fast_shuffle_is_the_reference_shuffle(G, O):
    """include/RansacLib/utils.h: the drivers' Mt19937 has std::mt19937's stream and their
    RandomShuffleAndResize (real swaps only for the steps that can reach the kept prefix, the other draws
    consumed block-wise with a vectorisable Lemire test) keeps the same prefix and leaves the engine where the
    reference's full Fisher-Yates would: golden vectors of the reference, then repeated shuffles on one
    engine (block boundaries, rejected draws at large n) against the oracle / the reference headers."""
    from ransaclib_b200 import host as H

    mt = np.random.RandomState(1234)  # numpy's legacy seeding is init_genrand, like std::mt19937::seed
    _, _, first = H.shuffle_and_resize(1234, 5, 2)
    assert first.tolist() == [int(x) for x in mt.randint(0, 2 ** 32, 8, dtype=np.uint64)]
    for c in G["shuffle_and_resize"]:
        out, nxt, _ = H.shuffle_and_resize(c["seed"], c["n"], c["target"])
        assert out.tolist() == c["out"] and nxt == c["next_draw"]
    ref = ol.reference() or O
    cases = [(1, 2, 1, 3), (2, 3, 2, 5), (3, 14, 14, 40), (4, 15, 14, 40), (5, 100, 6, 30), (6, 623, 14, 7),
             (7, 624, 14, 7), (8, 625, 14, 7), (9, 5000, 14, 50), (10, 5000, 28, 50), (11, 100000, 14, 12),
             (12, 2000000, 6, 3), (13, 40, 32, 9), (14, 10, 14, 4), (15, 1, 1, 2), (16, 0, 0, 2), (17, 300, 100, 5),
             (18, 65, 64, 9), (19, 66, 65, 9)]
    for seed, n, target, reps in cases:
        want = np.empty(max(target, 1), np.int32)
        wn = C.c_uint32()
        ref.shuffle_and_resize_reps(seed, n, target, reps, want, C.byref(wn))
        out, nxt, _ = H.shuffle_and_resize(seed, n, target, reps)
        assert out.tolist() == want[:target].tolist(), (seed, n, target, reps)
        assert nxt == wn.value, (seed, n, target, reps)
        if target <= n:  # the view-based variant the drivers run on a fused pass's pinned list
            got, nxt2 = H.shuffle_from_list(seed, n, target, reps)
            assert got.tolist() == (3 * want[:target] + 1).tolist() and nxt2 == wn.value, (seed, n, target, reps)


def test_pose_scoring_and_inliers(G, O):
    p = G["pose"]
    p2, p3, models = np.array(p["points2D"]), np.array(p["points3D"]), np.array(p["models"])
    s, c = ol.pose_score(O, p["fx"], p["fy"], p2, p3, models, p["sq_thr"])
    assert s.tolist() == p["scores"]  # same compiler flags, same operation order: bit-exact
    assert c.tolist() == p["counts"]
    assert c[-1] == 0 and s[-1] == p["sq_thr"] * len(p2)  # behind the camera: all outliers
    for m, want in zip(models[:3], p["inliers_first3"]):
        assert ol.pose_inliers(O, p["fx"], p["fy"], p2, p3, m, p["sq_thr"]).tolist() == want
    err = np.empty(len(p2))
    O.pose_errors(ol.c64(models[1]), p["fx"], p["fy"], len(p2), ol.c64(p2), ol.c64(p3), err)
    assert err.tolist() == p["errors_model1"]


def test_pose_lomsac_end_to_end(G, O):
    p = G["pose"]
    p2, p3 = np.array(p["points2D"]), np.array(p["points3D"])
    for c in p["lomsac"]:
        o = ol.Options.default(sq_thr=p["sq_thr"], seed=c["seed"], final_lsq=c["final_lsq"], lo_steps=c["lo_steps"],
                               min_iters=c["min_iters"], max_iters=c["max_iters"])
        st = ol.Stats()
        model = np.zeros(12)
        inl = np.zeros(len(p2), np.int32)
        O.lomsac_pose(C.byref(o), p["fx"], p["fy"], len(p2), ol.c64(p2), ol.c64(p3), model, C.byref(st), inl)
        assert st.num_iterations == c["num_iterations"]
        assert st.best_num_inliers == c["best_num_inliers"]
        assert st.number_lo_iterations == c["number_lo_iterations"]
        assert st.best_model_score == c["best_model_score"]
        assert st.inlier_ratio == c["inlier_ratio"]
        assert model.tolist() == c["model"]
        assert inl[:st.best_num_inliers].tolist() == c["inliers"]


def test_line_paths(G, O):
    l = G["line"]
    xy = np.array(l["xy"])
    smp = np.asarray(l["samples"], np.int32)
    m, v = np.empty((len(smp), 3)), np.empty(len(smp), np.uint8)
    O.line_minimal_batch(len(xy), ol.c64(xy), len(smp), smp, m, v)
    assert m.tolist() == l["models"] and v.tolist() == l["valid"]
    s, c = np.empty(len(m)), np.empty(len(m), np.int32)
    O.line_score_models(len(xy), ol.c64(xy), len(m), m, l["sq_thr"], s, c)
    assert s.tolist() == l["scores"] and c.tolist() == l["counts"]
    idx = np.empty(len(xy), np.int32)
    n = O.line_inliers(len(xy), ol.c64(xy), ol.c64(m[l["best"]]), l["sq_thr"], idx)
    assert idx[:n].tolist() == l["inliers_best"]
    for c in l["lomsac"]:
        o = ol.Options.default(sq_thr=l["sq_thr"], seed=c["seed"], final_lsq=c["final_lsq"], min_iters=100,
                               max_iters=100000)
        st = ol.Stats()
        model = np.zeros(3)
        inl = np.zeros(len(xy), np.int32)
        O.lomsac_line(C.byref(o), len(xy), ol.c64(xy), model, C.byref(st), inl)
        assert (st.num_iterations, st.best_num_inliers, st.number_lo_iterations) == (
            c["num_iterations"], c["best_num_inliers"], c["number_lo_iterations"])
        assert st.best_model_score == c["best_model_score"]
        assert model.tolist() == c["model"]
        assert inl[:st.best_num_inliers].tolist() == c["inliers"]


def test_hybrid_paths(G, O):
    h = G["hybrid"]
    xy0, p1, hm = np.array(h["xy0"]), np.array(h["xynn1"]), np.array(h["models"])
    s1 = np.asarray(h["samples1"], np.int32)
    m1, v1 = np.empty((len(s1), 3)), np.empty(len(s1), np.uint8)
    O.point_normal_batch(len(p1), ol.c64(p1), len(s1), s1, m1, v1)
    assert m1.tolist() == h["models"][6:]
    s, c = np.empty(len(hm)), np.empty(len(hm), np.int32)
    O.hybrid_line_score_models(len(xy0), ol.c64(xy0), len(p1), ol.c64(p1), len(hm), ol.c64(hm),
                               np.array(h["sq_thr"]), np.array(h["weights"]), s, c)
    assert s.tolist() == h["scores"] and c.tolist() == h["counts"]
    i0, i1, cc = np.empty(len(xy0), np.int32), np.empty(len(p1), np.int32), np.empty(2, np.int32)
    O.hybrid_line_inliers(len(xy0), ol.c64(xy0), len(p1), ol.c64(p1), ol.c64(hm[h["best"]]), np.array(h["sq_thr"]),
                          i0, i1, cc)
    assert i0[:cc[0]].tolist() == h["inliers0"] and i1[:cc[1]].tolist() == h["inliers1"]
    w0, w1 = np.array(h["weights0"]), np.array(h["weights1"])
    for key in ("lomsac", "lomsac_biased"):
        for c in h[key]:
            o = ol.HybridOptions.default(sq_thr=[1e-4, 1e-4], weights=[2.0, 0.5], seed=c["seed"],
                                         final_lsq=c["final_lsq"], min_iters=100, max_iters=10000,
                                         max_iters_per_solver=1000)
            st = ol.HybridStats()
            model = np.zeros(3)
            a, b = np.zeros(len(xy0), np.int32), np.zeros(len(p1), np.int32)
            if key == "lomsac":
                O.hybrid_lomsac_line(C.byref(o), len(xy0), ol.c64(xy0), len(p1), ol.c64(p1), np.array([0.2, 0.8]),
                                     model, C.byref(st), a, b)
            else:  # HybridBiasedSampling as the Sampler (hybrid_sampling.h:146-223)
                O.hybrid_lomsac_line_biased(C.byref(o), len(xy0), ol.c64(xy0), w0, len(p1), ol.c64(p1), w1,
                                            np.array([0.2, 0.8]), model, C.byref(st), a, b)
            assert st.num_iterations_total == c["num_iterations_total"]
            assert list(st.num_iterations_per_solver) == c["num_iterations_per_solver"]
            assert st.best_num_inliers == c["best_num_inliers"]
            assert st.best_solver_type == c["best_solver_type"]
            assert st.best_model_score == c["best_model_score"]
            assert list(st.inlier_ratios) == c["inlier_ratios"]
            assert st.number_lo_iterations == c["number_lo_iterations"]
            assert model.tolist() == c["model"]
            assert a[:st.num_inliers_per_type[0]].tolist() == c["inliers0"]
            assert b[:st.num_inliers_per_type[1]].tolist() == c["inliers1"]
    assert h["lomsac"][0]["model"] != h["lomsac_biased"][0]["model"]  # the sampler does change the run


def test_p3p_against_opencv(O):
    with open(os.path.join(HERE, "golden", "p3p_cv2.json")) as fh:
        cases = json.load(fh)["cases"]
    for c in cases:
        y, X = np.array(c["y"]), np.array(c["X"])
        Rs, ts = np.zeros(36), np.zeros(12)
        n = O.p3p(y, X, Rs, ts)
        ours = [(Rs[9 * i:9 * i + 9], ts[3 * i:3 * i + 3]) for i in range(n)]
        # ground truth among ours
        assert min(np.abs(R - np.array(c["R_gt"])).max() + np.abs(t - np.array(c["t_gt"])).max()
                   for R, t in ours) < 1e-7
        # same solution set as OpenCV (positive-depth solutions)
        assert n == len(c["cv2_solutions"])
        for s in c["cv2_solutions"]:
            assert min(np.abs(R - np.array(s["R"])).max() + np.abs(t - np.array(s["t"])).max()
                       for R, t in ours) < 1e-5
        # invariants: proper rotations that reproduce the bearings
        for R, t in ours:
            R = R.reshape(3, 3)
            assert abs(np.linalg.det(R) - 1) < 1e-8 and np.abs(R @ R.T - np.eye(3)).max() < 1e-8
            P = X.reshape(3, 3) @ R.T + t
            assert (P[:, 2] > 0).all()
            assert np.abs(P / np.linalg.norm(P, axis=1, keepdims=True) - y.reshape(3, 3)).max() < 1e-8


def test_reference_build_agrees_when_present(G, O):
    """In the build container oracle/_ref exists: oracle == reference on fresh random input."""
    ref = ol.reference()
    if ref is None:
        pytest.skip("oracle/_ref not built here")
    from ransaclib_b200 import synth

    inst = synth.pose_instance(77, 300, 300)
    f = inst["focal"]
    o = ol.Options.default(sq_thr=144.0, seed=5, final_lsq=1, min_iters=100, max_iters=5000)
    res = []
    for lib in (O, ref):
        st = ol.Stats()
        model = np.zeros(12)
        inl = np.zeros(600, np.int32)
        lib.lomsac_pose(C.byref(o), f, f, 600, inst["points2D"], inst["points3D"], model, C.byref(st), inl)
        res.append((st.num_iterations, st.best_num_inliers, st.best_model_score, model.tolist(),
                    inl[:st.best_num_inliers].tolist()))
    assert res[0] == res[1]


def test_user_defined_solver_oracle_equals_reference_headers(O):
    """The circle solver (not a reference example) through the restated LO-MSAC == through the reference's own
    unmodified ransac.h; this is the CPU side of tests/test_gpu_end_to_end.py::test_user_defined_solver_circle."""
    ref = ol.reference()
    if ref is None:
        pytest.skip("oracle/_ref not built here")
    rng = np.random.default_rng(12)
    n = 4000
    ang = rng.uniform(0, 2 * np.pi, n)
    xy = np.stack([1.5 + 2.0 * np.cos(ang), -0.5 + 2.0 * np.sin(ang)], 1) + rng.normal(0, 0.01, (n, 2))
    xy[::3] = rng.uniform(-3, 4, (len(xy[::3]), 2))
    xy = np.ascontiguousarray(xy)
    for seed, final_lsq in ((1, 0), (2, 1)):
        o = ol.Options.default(sq_thr=0.03 ** 2, seed=seed, final_lsq=final_lsq, min_iters=100, max_iters=5000)
        res = []
        for lib in (O, ref):
            st, model, inl = ol.Stats(), np.zeros(3), np.zeros(n, np.int32)
            lib.lomsac_circle(C.byref(o), n, xy, model, C.byref(st), inl)
            res.append((st.num_iterations, st.number_lo_iterations, st.best_num_inliers, st.best_model_score,
                        model.tolist(), inl[:st.best_num_inliers].tolist()))
        assert res[0] == res[1]
        assert res[0][2] > 2000 and abs(res[0][4][2] - 2.0) < 0.02


def test_hybrid_pose_solver_oracle_equals_reference_headers(O):
    """HybridPoseSolver (2D-3D + 3D-3D matches; the use of hybrid RANSAC the reference's README names without an
    example) through the restated hybrid LO-MSAC == through the reference's own unmodified hybrid_ransac.h, and the
    pose it finds is the ground truth. CPU side of tests/test_gpu_end_to_end.py::test_hybrid_pose_estimation."""
    from ransaclib_b200 import synth

    ref = ol.reference()
    inst = synth.hybrid_pose_instance(31, 250, 250, 150, 150)
    f = inst["focal"]
    for seed, priors, final_lsq in ((1, (0.5, 0.5), 0), (2, (0.9, 0.1), 1), (3, (0.1, 0.9), 0)):
        o = ol.HybridOptions.default(sq_thr=(144.0, 0.05 ** 2), weights=(1.0, 1.0), seed=seed, min_iters=100,
                                     max_iters=5000, max_iters_per_solver=5000, final_lsq=final_lsq)
        res = []
        for lib in (O, ref):
            if lib is None:
                continue
            m, st, a, b = ol.hybrid_lomsac_pose(lib, o, f, f, inst["points2D"], inst["points3D"], inst["cam1"],
                                                inst["world1"], priors)
            res.append((st.num_iterations_total, list(st.num_iterations_per_solver), st.best_solver_type,
                        st.number_lo_iterations, st.best_model_score, m.tolist(), a.tolist(), b.tolist()))
        if len(res) == 2:
            assert res[0] == res[1]
        m = np.array(res[0][5])
        R = m[:9].reshape(3, 3).T  # column-major
        assert np.abs(R - inst["R_gt"]).max() < 5e-3 and np.abs(m[9:] - inst["c_gt"]).max() < 5e-2
        assert len(res[0][6]) >= 240 and len(res[0][7]) >= 140
        # the inlier lists are what EvaluateModelOnPoint says about the returned model
        e0, e1 = np.zeros(500), np.zeros(300)
        O.hybrid_pose_errors(m, f, f, 500, inst["points2D"], inst["points3D"], 300, inst["cam1"], inst["world1"], e0, e1)
        assert np.flatnonzero(e0 < 144.0).tolist() == res[0][6]
        assert np.flatnonzero(e1 < 0.05 ** 2).tolist() == res[0][7]
