"""End-to-end LO-MSAC wall time of BASELINE configs[0..2] (C1 line, C2 pose with 10k forced
hypotheses, C3 hybrid line N0 = N1 = 100k) through the drop-in template layer on one GPU, next to the reference's own
EstimateModel (oracle/_ref/libransac_ref_fast.so: reference headers, reference flags) on one host core, same data and
seeds. Checks on the way that both runs agree (iterations, inlier count). Prints one JSON object.

    python tests/e2e_configs.py > gpurun_out/<tag>_e2e_configs.json
"""
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import oracle_lib as ol  # noqa: E402  (lives under tests/: the oracle is test infrastructure)
from ransaclib_b200 import host as H  # noqa: E402
from ransaclib_b200 import synth  # noqa: E402


def best_of(fn, reps=3):
    out, best = None, 1e30
    for _ in range(reps):
        t0 = time.perf_counter()
        out = fn()
        best = min(best, time.perf_counter() - t0)
    return out, best


def main():
    ref = ol.reference(fast=True) or ol.oracle()
    kind = "reference" if ol.reference(fast=True) is not None else "port"
    rows = []

    # ---- C1: line_estimation.cc, N = 10,000, thr^2 = 1e-4, max 100k iterations ----
    for ratio in (0.1, 0.5, 0.9, 0.99):
        n, seed = 10000, 42
        n_out = int(n * ratio)
        xy = synth.line_instance(1042, n - n_out, n_out)
        o = ol.Options.default(sq_thr=1e-4, seed=seed, min_iters=100, max_iters=100000)

        def cpu():
            st, m, inl = ol.Stats(), np.zeros(3), np.zeros(n, np.int32)
            ref.lomsac_line(C.byref(o), n, ol.c64(xy), m, C.byref(st), inl)
            return st

        opts = H.LORansacOptions(squared_inlier_threshold_=1e-4, random_seed_=seed, min_num_iterations_=100,
                                 max_num_iterations_=100000)
        ost, t_cpu = best_of(cpu)
        row = {"config": f"C1 line N=10000 outlier_ratio={ratio}", "iterations": ost.num_iterations,
               "inliers": ost.best_num_inliers, "cpu_s": t_cpu}
        # host-driven loop (one launch per pass), f64 and f32 sweeps
        os.environ["RLB_HOST_DRIVER"] = "1"
        for name, prec in (("gpu_f64_s", 0), ("gpu_f32_s", 1)):
            (m, st), t = best_of(lambda: H.estimate_line(opts, xy, precision=prec))
            assert st["num_iterations"] == ost.num_iterations and st["best_num_inliers"] == ost.best_num_inliers
            row[name] = t
            row["single_passes"], row["batches"] = st["single_passes"], st["batches"]
        os.environ.pop("RLB_HOST_DRIVER")
        # device-resident driver (the default of the drop-in layer for LineEstimator): the whole run in one launch;
        # wall time includes the upload of the data set at solver construction, like the rows above
        (m, st), t = best_of(lambda: H.estimate_line(opts, xy), reps=5)
        assert st["num_iterations"] == ost.num_iterations and st["best_num_inliers"] == ost.best_num_inliers
        row["handed_back_to_host_loop"] = bool(st["batches"] > 0)  # runs needing > 3000 hypotheses after 384
        row["gpu_device_resident_s"] = t
        row["device_resident_estimate_s"] = st["seconds"]  # EstimateModel alone (solver already constructed)
        row["device_resident_passes"] = st["single_passes"]
        row["speedup_device_resident"] = t_cpu / t
        rows.append(row)

    # ---- C2: camera_pose_estimation.cc, N = 10k, 50 % outliers, min = max = 10,000 hypotheses ----
    for forced in (True, False):
        inst = synth.pose_instance(4242, 5000, 5000)
        f, p2, p3 = inst["focal"], inst["points2D"], inst["points3D"]
        mn = 10000 if forced else 100
        o = ol.Options.default(sq_thr=144.0, seed=42, min_iters=mn, max_iters=10000)

        def cpu():
            st, m, inl = ol.Stats(), np.zeros(12), np.zeros(len(p2), np.int32)
            ref.lomsac_pose(C.byref(o), f, f, len(p2), ol.c64(p2), ol.c64(p3), m, C.byref(st), inl)
            return st

        opts = H.LORansacOptions(squared_inlier_threshold_=144.0, random_seed_=42, min_num_iterations_=mn,
                                 max_num_iterations_=10000)
        ost, t_cpu = best_of(cpu, reps=2 if forced else 3)
        row = {"config": "C2 pose N=10000, " + ("10k hypotheses forced" if forced else "adaptive stop"),
               "iterations": ost.num_iterations, "inliers": ost.best_num_inliers, "cpu_s": t_cpu}
        for name, prec in (("gpu_f64_s", 0), ("gpu_f32_s", 1)):
            (m, st), t = best_of(lambda: H.estimate_pose(opts, f, f, p2, p3, precision=prec))
            assert st["num_iterations"] == ost.num_iterations and st["best_num_inliers"] == ost.best_num_inliers
            row[name] = t
            row["single_passes"], row["batches"] = st["single_passes"], st["batches"]
        rows.append(row)

    # ---- C3: hybrid_line_estimation.cc, N0 = N1 = 100k, 50 % outliers ----
    n = 100000
    xy0, p1 = synth.hybrid_line_instance(306, n // 2, n // 2, n // 2, n // 2)
    ho = ol.HybridOptions.default(sq_thr=[1e-4, 1e-4], weights=[2.0, 0.5], seed=6, min_iters=100, max_iters=10000,
                                  max_iters_per_solver=1000)

    def cpu_h():
        st, m = ol.HybridStats(), np.zeros(3)
        a, b = np.zeros(n, np.int32), np.zeros(n, np.int32)
        ref.hybrid_lomsac_line(C.byref(ho), n, ol.c64(xy0), n, ol.c64(p1), np.array([0.2, 0.8]), m, C.byref(st), a, b)
        return st

    hopts = H.HybridLORansacOptions(squared_inlier_thresholds_=[1e-4, 1e-4], data_type_weights_=[2.0, 0.5],
                                    random_seed_=6, min_num_iterations_=100, max_num_iterations_=10000,
                                    max_num_iterations_per_solver_=1000)
    ost, t_cpu = best_of(cpu_h)
    row = {"config": "C3 hybrid line N0=N1=100000", "iterations": ost.num_iterations_total,
           "inliers": ost.best_num_inliers, "cpu_s": t_cpu}
    for name, prec in (("gpu_f64_s", 0), ("gpu_f32_s", 1)):
        (m, st), t = best_of(lambda: H.estimate_hybrid_line(hopts, xy0, p1, [0.2, 0.8], precision=prec))
        assert st["num_iterations_total"] == ost.num_iterations_total and st["best_num_inliers"] == ost.best_num_inliers
        row[name] = t
    rows.append(row)

    # ---- a hybrid POSE solver (examples/hybrid_pose_estimator.h; no reference example, the reference's own
    #      hybrid_ransac.h over the oracle's counterpart is the CPU side): 50k 2D-3D + 20k 3D-3D matches ----
    hp = synth.hybrid_pose_instance(33, 25000, 25000, 10000, 10000)
    fh = hp["focal"]
    thr = (144.0, 0.05 ** 2)
    po = ol.HybridOptions.default(sq_thr=thr, weights=(1.0, 1.0), seed=7, min_iters=100, max_iters=10000,
                                  max_iters_per_solver=10000)
    popts = H.HybridLORansacOptions(squared_inlier_thresholds_=list(thr), data_type_weights_=[1.0, 1.0],
                                    random_seed_=7, min_num_iterations_=100, max_num_iterations_=10000,
                                    max_num_iterations_per_solver_=10000)
    (om, ost, a, b), t_cpu = best_of(lambda: ol.hybrid_lomsac_pose(ref, po, fh, fh, hp["points2D"], hp["points3D"],
                                                                   hp["cam1"], hp["world1"], (0.5, 0.5)))
    row = {"config": "hybrid pose, 50000 2D-3D + 20000 3D-3D matches (user-defined HybridSolver)",
           "iterations": ost.num_iterations_total, "inliers": ost.best_num_inliers, "cpu_s": t_cpu}
    for name, prec in (("gpu_f64_s", 0), ("gpu_f32_s", 1)):
        (m, st), t = best_of(lambda: H.estimate_hybrid_pose(popts, fh, fh, hp["points2D"], hp["points3D"], hp["cam1"],
                                                            hp["world1"], (0.5, 0.5), precision=prec))
        assert st["num_iterations_total"] == ost.num_iterations_total and st["best_num_inliers"] == ost.best_num_inliers
        assert st["inlier_indices"][0].tolist() == a.tolist() and st["inlier_indices"][1].tolist() == b.tolist()
        row[name] = t
    rows.append(row)

    for r in rows:
        r["speedup_f64"] = r["cpu_s"] / r["gpu_f64_s"]
        r["speedup_f32"] = r["cpu_s"] / r["gpu_f32_s"]
    print(json.dumps({"what": "end-to-end EstimateModel wall time, best of 3; GPU = drop-in template layer incl. the "
                              "upload of the data set at solver construction; CPU = " + kind + " on one core",
                      "cpu_kind": kind, "rows": rows}, indent=1))


if __name__ == "__main__":
    main()
