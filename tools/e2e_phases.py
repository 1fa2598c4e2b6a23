"""Times the phases of bench.py's e2e step (host-buffer C ABI path) on one GPU."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import ransaclib_b200 as rlb
import bench

n, k = 1_000_000, int(sys.argv[1]) if len(sys.argv) > 1 else 65536
p2, p3, f, inst = bench.make_instance(n)
samples = bench.make_samples(n, k)
pin = lambda a: torch.from_numpy(a).pin_memory().numpy()
h_p2, h_p3, h_s = pin(p2), pin(p3), pin(samples)
ctx = rlb.Context(0)
for it in range(4):
    t = [time.perf_counter()]
    sc = ctx.pose_scorer(h_p2, h_p3, f, f); t.append(time.perf_counter())
    r = sc.hypothesize_and_score(h_s, 144.0, precision=rlb.RLB_F32, want_models=False); t.append(time.perf_counter())
    lists, counts, score = sc.get_inliers(r["best_model"], 144.0, 144.0); t.append(time.perf_counter())
    sc.close(); t.append(time.perf_counter())
    print("iter", it, "create %.2f ms | hyp+score %.2f ms | inliers %.2f ms | close %.2f ms" % tuple(
        1e3 * (b - a) for a, b in zip(t, t[1:])))
