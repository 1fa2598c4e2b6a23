"""Device-resident line LO-MSAC (one CTA) against the host-driven loop as the problem grows: where does the
one-launch driver stop paying? Prints one line per size (best of 3, wall clock, data set resident)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import ransaclib_b200 as rlb
from ransaclib_b200 import host as H, synth

ctx = rlb.Context(0)
for n in (2000, 10000, 16000, 20000, 24000, 30000, 60000, 120000, 250000):
    xy = synth.line_instance(77, n // 2, n - n // 2)
    kw = dict(sq_thr=1e-4, min_iters=100, max_iters=100000)
    sc = ctx.line_scorer(xy)
    opt = ctx.lomsac_options(**kw)
    best_dev = 1e9
    for _ in range(4):
        t0 = time.perf_counter(); line, r, inl = sc.line_lomsac(opt, 42); best_dev = min(best_dev, time.perf_counter() - t0)
    sc.close()
    opts = H.LORansacOptions(squared_inlier_threshold_=1e-4, random_seed_=42, min_num_iterations_=100, max_num_iterations_=100000)
    os.environ["RLB_HOST_DRIVER"] = "1"
    best_host = 1e9
    for _ in range(4):
        m, st = H.estimate_line(opts, xy); best_host = min(best_host, st["seconds"])
    os.environ.pop("RLB_HOST_DRIVER")
    assert st["num_iterations"] == r["num_iterations"] and st["best_num_inliers"] == r["num_inliers"]
    print(f"n={n:7d} iterations={r['num_iterations']} device-resident {best_dev*1e3:7.2f} ms   host-driven {best_host*1e3:7.2f} ms")
