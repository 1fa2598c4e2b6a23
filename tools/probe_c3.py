"""configs[2] (hybrid line, N0 = N1 = 100k): where the end-to-end time goes. One fused K=1 pass timed through the C ABI
with and without the inlier list, scorer construction, and the whole estimate. Prints one JSON object."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ransaclib_b200 as rlb  # noqa: E402
from ransaclib_b200 import host as H, synth  # noqa: E402


def best(f, reps=30):
    f()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        f()
        ts.append(time.perf_counter() - t0)
    return float(np.median(ts)) * 1e6


def main():
    n = 100000
    xy0, p1 = synth.hybrid_line_instance(306, n // 2, n // 2, n // 2, n // 2)
    hopts = H.HybridLORansacOptions(squared_inlier_thresholds_=[1e-4, 1e-4], data_type_weights_=[2.0, 0.5],
                                    random_seed_=6, min_num_iterations_=100, max_num_iterations_=10000,
                                    max_num_iterations_per_solver_=1000)
    (m, st) = H.estimate_hybrid_line(hopts, xy0, p1, [0.2, 0.8], precision=0)
    out = {"estimate_us": best(lambda: H.estimate_hybrid_line(hopts, xy0, p1, [0.2, 0.8], precision=0), 10),
           "stats": {k: v for k, v in st.items() if np.isscalar(v)}}
    ctx = rlb.Context(0)
    out["scorer_create_close_us"] = best(lambda: ctx.hybrid_line_scorer(xy0, p1).close(), 10)
    s = ctx.hybrid_line_scorer(xy0, p1)
    thr, w = [1e-4, 1e-4], [2.0, 0.5]
    out["pass_with_list_us"] = best(lambda: s.get_inliers(m, thr, thr, w, want_indices=True))
    out["pass_count_only_us"] = best(lambda: s.get_inliers(m, thr, thr, w, want_indices=False))
    bad = np.array([1.0, 0.0, 5.0])
    out["pass_with_empty_list_us"] = best(lambda: s.get_inliers(bad, thr, thr, w, want_indices=True))
    print(json.dumps(out))


if __name__ == "__main__":
    main()
