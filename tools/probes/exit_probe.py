"""Does a script that leaves a context and a scorer alive at interpreter exit terminate cleanly? Runs the case in a
child process (faulthandler on) and reports its return code; variants: timing on/off, with/without get_inliers."""
import subprocess
import sys

CASE = r'''
import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np
import ransaclib_b200 as rlb
from ransaclib_b200 import synth
from ransaclib_b200.sampling import UniformSampling
variant = sys.argv[1]
inst = synth.pose_instance(4242, 5000, 5000)
f, p2, p3 = inst["focal"], inst["points2D"], inst["points3D"]
ctx = rlb.Context(0)
sc = ctx.pose_scorer(p2, p3, f, f)
samples = UniformSampling(42, len(p2), 4).sample_fast(256)
r = sc.hypothesize_and_score(samples, 144.0)
m = r["best_model"]
if "timing" in variant:
    ctx.set_timing(True)
if "inl" in variant:
    sc.get_inliers(m, 144.0, 144.0)
if "ktime" in variant:
    print(ctx.kernel_time_ms(1))
if "lambda" in variant:
    fn = lambda: sc.get_inliers(m, 144.0, 144.0)
    fn()
print("body done", variant, flush=True)
'''

for variant in ("plain", "inl", "timing_inl", "timing_inl_ktime", "timing_inl_ktime_lambda"):
    r = subprocess.run([sys.executable, "-X", "faulthandler", "-c", CASE, variant], capture_output=True, text=True)
    print("==", variant, "rc", r.returncode)
    print(r.stdout[-300:])
    if r.returncode != 0:
        print(r.stderr[-2500:])
