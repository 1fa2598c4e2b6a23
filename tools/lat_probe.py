import sys, os, time
sys.path.insert(0, os.getcwd())
import numpy as np
import ransaclib_b200 as rlb
from ransaclib_b200 import synth
from ransaclib_b200.sampling import UniformSampling
inst = synth.pose_instance(4242, 5000, 5000)
f, p2, p3 = inst["focal"], inst["points2D"], inst["points3D"]
ctx = rlb.Context(0)
def t(fn, reps):
    fn(); t0 = time.perf_counter()
    for _ in range(reps): r = fn()
    return (time.perf_counter() - t0) / reps * 1e6
def create():
    s = ctx.pose_scorer(p2, p3, f, f); s.close()
print("scorer create+destroy us", round(t(create, 50), 1))
sc = ctx.pose_scorer(p2, p3, f, f)
samples = UniformSampling(42, len(p2), 4).sample_fast(4096)
r = sc.hypothesize_and_score(samples[:256], 144.0)
m = r["best_model"]
print("pass with list us", round(t(lambda: sc.get_inliers(m, 144.0, 144.0), 300), 1))
print("pass count only us", round(t(lambda: sc.get_inliers(m, 144.0, 144.0, want_indices=False), 300), 1))
for k in (256, 512, 1024, 2048, 4096):
    print("hyp+score K=%d us" % k, round(t(lambda: sc.hypothesize_and_score(samples[:k], 144.0), 50), 1),
          "f32", round(t(lambda: sc.hypothesize_and_score(samples[:k], 144.0, precision=rlb.RLB_F32), 50), 1))
ctx.set_timing(True)
ks = []
for _ in range(50):
    sc.get_inliers(m, 144.0, 144.0)
    ks.append(ctx.kernel_time_ms(1) * 1e3)
print("pass kernel (events around the launch) us", round(float(np.median(ks)), 1))
print("pass with list, timing events on, us", round(t(lambda: sc.get_inliers(m, 144.0, 144.0), 300), 1))
