"""configs[1] (one pose problem, N = 10k) through the device-resident batched-localization kernel as a batch of ONE
query, against the host-driven drop-in loop: would a DeviceResident route for CalibratedAbsolutePoseEstimator pay?"""
import os, sys, time
sys.path.insert(0, os.getcwd())
import numpy as np
import ransaclib_b200 as rlb
from ransaclib_b200 import host as H, synth

inst = synth.pose_instance(4242, 5000, 5000)
f, p2, p3 = inst["focal"], inst["points2D"], inst["points3D"]
ctx = rlb.Context(0)
o = ctx.localize_options(inlier_threshold=12.0, number_lo_steps=10, min_num_iterations=100, max_num_iterations=10000) \
    if hasattr(ctx, "localize_options") else None
def best(fn, reps=10):
    fn(); ts = []
    for _ in range(reps):
        t0 = time.perf_counter(); r = fn(); ts.append(time.perf_counter() - t0)
    return r, min(ts) * 1e3
off = np.array([0, len(p2)], np.int64)
(poses, res, inl), t_dev = best(lambda: ctx.localize_batch(off, p2, p3, f, f, [7], o))
print("device-resident, one query of %d matches: %.3f ms" % (len(p2), t_dev), res)
opts = H.LORansacOptions(squared_inlier_threshold_=144.0, random_seed_=7, min_num_iterations_=100, max_num_iterations_=10000)
(m, st), t_host = best(lambda: H.estimate_pose(opts, f, f, p2, p3))
print("host-driven drop-in loop: %.3f ms" % t_host, {k: st[k] for k in ("num_iterations", "best_num_inliers", "number_lo_iterations", "single_passes")})
