"""Generates tests/golden/reference_vectors.json by running the REFERENCE's own
unmodified headers (oracle/_ref/libransac_ref.so, built from /root/reference by
oracle/Makefile) on small seeded inputs. Run in the build container:

    make -C oracle && python tests/golden/make_golden.py

The inputs are stored in the file too, so the fixtures do not depend on numpy's
generator staying stable. Everything is exact: Python's repr round-trips doubles.
"""
import ctypes as C
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import oracle_lib as ol  # noqa: E402
from ransaclib_b200 import synth  # noqa: E402


def main():
    ref = ol.reference()
    assert ref is not None, "build oracle/_ref first (needs /root/reference)"
    out = {"generator": "tests/golden/make_golden.py over oracle/_ref/libransac_ref.so "
                        "(reference RansacLib/*.h @ b952083f, g++ 13.3 -O2 -ffp-contract=off)"}

    # ---- samplers / utils (sampling.h, hybrid_sampling.h, utils.h) ----------
    smp = []
    for seed, n, s, k in [(0, 10000, 4, 8), (42, 10000, 4, 8), (42, 10000, 2, 8), (42, 5, 4, 6), (7, 6, 2, 6),
                          (3, 1000000, 4, 4)]:
        smp.append({"seed": seed, "n": n, "s": s, "k": k, "samples": ol.sample_uniform(ref, seed, n, s, k).tolist()})
    out["uniform_sampling"] = smp
    hyb = []
    for seed, nd, sizes in [(42, [100000, 100000], [[2, 0], [0, 1], [2, 0], [0, 1], [2, 0]]),
                            (5, [3, 4], [[2, 0], [0, 1], [2, 0], [0, 1]])]:
        k = len(sizes)
        o = np.empty((k, 2), np.int32)
        ref.sample_hybrid(seed, 2, np.asarray(nd, np.int32), k, np.asarray(sizes, np.int32).reshape(-1), 2,
                          o.reshape(-1))
        hyb.append({"seed": seed, "num_data": nd, "sizes": sizes, "samples": o.tolist()})
    out["hybrid_sampling"] = hyb
    # HybridBiasedSampling (hybrid_sampling.h:146-223): weights are small integers / 4 (exact), some zero
    hb = []
    for seed, nd, sizes in [(42, [50, 40], [[2, 0], [0, 1], [2, 1], [3, 2], [0, 1], [2, 0]]),
                            (7, [6, 5], [[3, 0], [0, 2], [3, 2], [1, 1]]),      # "all candidates" short-cut
                            (1, [1000, 2000], [[2, 0], [0, 1]] * 4)]:
        w = [[((j * 7 + t * 3 + seed) % 5) * 0.25 for j in range(nd[t])] for t in range(2)]
        if seed == 7:
            w = [[1.0, 0.0, 2.0, 0.0, 0.5, 0.0], [0.0, 3.0, 0.0, 1.0, 0.0]]   # 3 and 2 candidates
        k = len(sizes)
        stride = max(sum(r) for r in sizes)
        o = np.empty((k, stride), np.int32)
        ref.sample_hybrid_biased(seed, 2, np.asarray(nd, np.int32), np.asarray(w[0] + w[1], np.float64), k,
                                 np.asarray(sizes, np.int32).reshape(-1), stride, o.reshape(-1))
        hb.append({"seed": seed, "num_data": nd, "weights": w, "sizes": sizes, "samples": o.tolist()})
    out["hybrid_biased_sampling"] = hb
    nri = []
    for args in [(0.5, 1e-4, 4, 100, 10000), (0.5, 1e-4, 2, 100, 100000), (0.1, 1e-4, 4, 100, 1000000),
                 (0.05, 1e-4, 4, 100, 1000000), (0.001, 1e-4, 2, 100, 100000), (0.0, 1e-4, 4, 100, 10000),
                 (1.0, 1e-4, 4, 100, 10000), (0.9, 1e-4, 4, 10, 10000), (0.37, 1e-4, 4, 100, 10000)]:
        nri.append({"args": list(args), "value": int(ref.num_required_iterations(*args))})
    out["num_required_iterations"] = nri
    nrh = []
    for ratios, sizes, mn, mx in [([0.5, 0.5], [2, 0], 100, 1000), ([0.5, 0.25], [0, 1], 100, 1000),
                                  ([0.1, 0.9], [2, 0], 10, 100000), ([0.0, 0.5], [2, 0], 100, 1000)]:
        v = ref.num_required_iterations_hybrid(2, np.asarray(ratios, np.float64), 1e-4, np.asarray(sizes, np.int32),
                                               mn, mx)
        nrh.append({"ratios": ratios, "sizes": sizes, "min": mn, "max": mx, "value": int(v)})
    out["num_required_iterations_hybrid"] = nrh
    shf = []
    for seed, n, target in [(42, 10, 4), (1, 100, 28), (9, 7, 7), (3, 2, 1)]:
        o = np.empty(target, np.int32)
        nxt = C.c_uint32()
        ref.shuffle_and_resize(seed, n, target, o, C.byref(nxt))
        shf.append({"seed": seed, "n": n, "target": target, "out": o.tolist(), "next_draw": int(nxt.value)})
    out["shuffle_and_resize"] = shf

    # ---- pose: ScoreModel / GetInliers / EstimateModel -----------------------
    inst = synth.pose_instance(11, 160, 140)
    f = inst["focal"]
    p2, p3 = inst["points2D"], inst["points3D"]
    gt = synth.pose_to_model(inst["R_gt"], inst["c_gt"])
    models = np.concatenate([gt[None], synth.perturbed_poses(5, inst["R_gt"], inst["c_gt"], 11)])
    flip = gt.copy()
    flip[:9] = -flip[:9]  # camera looking away: every point behind the camera
    models = np.concatenate([models, flip[None]])
    thr = 144.0
    scores, counts = ol.pose_score(ref, f, f, p2, p3, models, thr)
    inl = [ol.pose_inliers(ref, f, f, p2, p3, m, thr).tolist() for m in models[:3]]
    err = np.empty(len(p2))
    ref.pose_errors(ol.c64(models[1]), f, f, len(p2), ol.c64(p2), ol.c64(p3), err)
    pose = {"fx": f, "fy": f, "points2D": p2.tolist(), "points3D": p3.tolist(), "models": models.tolist(),
            "sq_thr": thr, "scores": scores.tolist(), "counts": counts.tolist(), "inliers_first3": inl,
            "errors_model1": err.tolist(), "lomsac": []}
    for seed, final_lsq, lo_steps in [(42, 0, 10), (7, 1, 10), (3, 0, 0)]:
        o = ol.Options.default(sq_thr=thr, seed=seed, final_lsq=final_lsq, lo_steps=lo_steps, min_iters=100,
                               max_iters=2000)
        st = ol.Stats()
        model = np.zeros(12)
        inliers = np.zeros(len(p2), np.int32)
        ref.lomsac_pose(C.byref(o), f, f, len(p2), ol.c64(p2), ol.c64(p3), model, C.byref(st), inliers)
        pose["lomsac"].append({"seed": seed, "final_lsq": final_lsq, "lo_steps": lo_steps, "min_iters": 100,
                               "max_iters": 2000, "num_iterations": st.num_iterations,
                               "best_num_inliers": st.best_num_inliers, "best_model_score": st.best_model_score,
                               "inlier_ratio": st.inlier_ratio, "number_lo_iterations": st.number_lo_iterations,
                               "model": model.tolist(), "inliers": inliers[:st.best_num_inliers].tolist()})
    out["pose"] = pose

    # ---- line ------------------------------------------------------------------
    xy = synth.line_instance(21, 180, 120)
    lmodels, lvalid = np.empty((12, 3)), np.empty(12, np.uint8)
    lsmp = ol.sample_uniform(ref, 5, len(xy), 2, 12)
    ref.line_minimal_batch(len(xy), ol.c64(xy), 12, lsmp, lmodels, lvalid)
    lthr = 1e-4
    ls, lc = np.empty(12), np.empty(12, np.int32)
    ref.line_score_models(len(xy), ol.c64(xy), 12, lmodels, lthr, ls, lc)
    li = np.empty(len(xy), np.int32)
    best = int(np.argmin(ls))
    c = ref.line_inliers(len(xy), ol.c64(xy), ol.c64(lmodels[best]), lthr, li)
    line = {"xy": xy.tolist(), "samples": lsmp.tolist(), "models": lmodels.tolist(), "valid": lvalid.tolist(),
            "sq_thr": lthr, "scores": ls.tolist(), "counts": lc.tolist(), "best": best,
            "inliers_best": li[:c].tolist(), "lomsac": []}
    for seed, final_lsq in [(42, 0), (8, 1)]:
        o = ol.Options.default(sq_thr=lthr, seed=seed, final_lsq=final_lsq, min_iters=100, max_iters=100000)
        st = ol.Stats()
        model = np.zeros(3)
        inliers = np.zeros(len(xy), np.int32)
        ref.lomsac_line(C.byref(o), len(xy), ol.c64(xy), model, C.byref(st), inliers)
        line["lomsac"].append({"seed": seed, "final_lsq": final_lsq, "num_iterations": st.num_iterations,
                               "best_num_inliers": st.best_num_inliers, "best_model_score": st.best_model_score,
                               "inlier_ratio": st.inlier_ratio, "number_lo_iterations": st.number_lo_iterations,
                               "model": model.tolist(), "inliers": inliers[:st.best_num_inliers].tolist()})
    out["line"] = line

    # ---- hybrid line -------------------------------------------------------------
    xy0, p1 = synth.hybrid_line_instance(31, 90, 70, 100, 60)
    hm = np.concatenate([lmodels[:0], np.empty((0, 3))])
    s0 = ol.sample_uniform(ref, 9, len(xy0), 2, 6)
    m0, v0 = np.empty((6, 3)), np.empty(6, np.uint8)
    ref.line_minimal_batch(len(xy0), ol.c64(xy0), 6, s0, m0, v0)
    s1 = ol.sample_uniform(ref, 10, len(p1), 1, 6).reshape(-1)
    s1 = np.ascontiguousarray(s1, np.int32)
    m1, v1 = np.empty((6, 3)), np.empty(6, np.uint8)
    ref.point_normal_batch(len(p1), ol.c64(p1), 6, s1, m1, v1)
    hm = np.concatenate([m0, m1])
    hthr = np.array([1e-4, 2e-4])
    hw = np.array([2.0, 0.5])
    hs, hc = np.empty(12), np.empty(12, np.int32)
    ref.hybrid_line_score_models(len(xy0), ol.c64(xy0), len(p1), ol.c64(p1), 12, ol.c64(hm), hthr, hw, hs, hc)
    hb = int(np.argmin(hs))
    i0, i1, cc = np.empty(len(xy0), np.int32), np.empty(len(p1), np.int32), np.empty(2, np.int32)
    ref.hybrid_line_inliers(len(xy0), ol.c64(xy0), len(p1), ol.c64(p1), ol.c64(hm[hb]), hthr, i0, i1, cc)
    hybrid = {"xy0": xy0.tolist(), "xynn1": p1.tolist(), "samples0": s0.tolist(), "samples1": s1.tolist(),
              "models": hm.tolist(), "valid": np.concatenate([v0, v1]).tolist(), "sq_thr": hthr.tolist(),
              "weights": hw.tolist(), "scores": hs.tolist(), "counts": hc.tolist(), "best": hb,
              "inliers0": i0[:cc[0]].tolist(), "inliers1": i1[:cc[1]].tolist(), "lomsac": []}
    for seed, final_lsq in [(42, 0), (6, 1)]:
        o = ol.HybridOptions.default(sq_thr=[1e-4, 1e-4], weights=[2.0, 0.5], seed=seed, final_lsq=final_lsq,
                                     min_iters=100, max_iters=10000, max_iters_per_solver=1000)
        st = ol.HybridStats()
        model = np.zeros(3)
        a, b = np.zeros(len(xy0), np.int32), np.zeros(len(p1), np.int32)
        ref.hybrid_lomsac_line(C.byref(o), len(xy0), ol.c64(xy0), len(p1), ol.c64(p1), np.array([0.2, 0.8]), model,
                               C.byref(st), a, b)
        hybrid["lomsac"].append({
            "seed": seed, "final_lsq": final_lsq, "num_iterations_total": st.num_iterations_total,
            "num_iterations_per_solver": list(st.num_iterations_per_solver),
            "best_num_inliers": st.best_num_inliers, "best_solver_type": st.best_solver_type,
            "best_model_score": st.best_model_score, "inlier_ratios": list(st.inlier_ratios),
            "number_lo_iterations": st.number_lo_iterations, "model": model.tolist(),
            "inliers0": a[:st.num_inliers_per_type[0]].tolist(), "inliers1": b[:st.num_inliers_per_type[1]].tolist()})
    # the same instance through HybridBiasedSampling: outliers of type 0 down-weighted, a third of type 1 never drawn
    rngw = np.random.default_rng(77)
    w0 = np.round(rngw.uniform(0.1, 1.0, len(xy0)), 3)
    w1 = np.round(rngw.uniform(0.1, 1.0, len(p1)), 3)
    w1[::3] = 0.0
    hybrid["weights0"], hybrid["weights1"], hybrid["lomsac_biased"] = w0.tolist(), w1.tolist(), []
    for seed, final_lsq in [(42, 0), (6, 1)]:
        o = ol.HybridOptions.default(sq_thr=[1e-4, 1e-4], weights=[2.0, 0.5], seed=seed, final_lsq=final_lsq,
                                     min_iters=100, max_iters=10000, max_iters_per_solver=1000)
        st = ol.HybridStats()
        model = np.zeros(3)
        a, b = np.zeros(len(xy0), np.int32), np.zeros(len(p1), np.int32)
        ref.hybrid_lomsac_line_biased(C.byref(o), len(xy0), ol.c64(xy0), ol.c64(w0), len(p1), ol.c64(p1), ol.c64(w1),
                                      np.array([0.2, 0.8]), model, C.byref(st), a, b)
        hybrid["lomsac_biased"].append({
            "seed": seed, "final_lsq": final_lsq, "num_iterations_total": st.num_iterations_total,
            "num_iterations_per_solver": list(st.num_iterations_per_solver),
            "best_num_inliers": st.best_num_inliers, "best_solver_type": st.best_solver_type,
            "best_model_score": st.best_model_score, "inlier_ratios": list(st.inlier_ratios),
            "number_lo_iterations": st.number_lo_iterations, "model": model.tolist(),
            "inliers0": a[:st.num_inliers_per_type[0]].tolist(), "inliers1": b[:st.num_inliers_per_type[1]].tolist()})
    out["hybrid"] = hybrid

    path = os.path.join(HERE, "reference_vectors.json")
    with open(path, "w") as fh:
        json.dump(out, fh)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
