"""Per-phase cycle counters of the localization kernel (debug_flags bit 1). usage: prof_localize.py Q"""
import sys

import numpy as np

sys.path.insert(0, ".")
import ransaclib_b200 as rlb
from ransaclib_b200 import synth

Q = int(sys.argv[1]) if len(sys.argv) > 1 else 1
p2s, p3s, fs = [], [], []
for q in range(Q):
    inst = synth.pose_instance(5000 + (q % 64), 1000, 1000)
    p2s.append(inst["points2D"]); p3s.append(inst["points3D"]); fs.append(float(np.float32(inst["focal"])))
p2, p3, f = np.concatenate(p2s), np.concatenate(p3s), np.array(fs)
off = np.arange(Q + 1, dtype=np.int64) * 2000
seeds = np.arange(42, 42 + Q, dtype=np.uint32)
ctx = rlb.Context(0)
o = ctx.localize_options(12.0, 10, 100, 10000)
o.debug_flags = 2
ctx.localize_batch(off, p2, p3, f, f, seeds, o)
poses, res, inl = ctx.localize_batch(off, p2, p3, f, f, seeds, o)
names = ["batches(sample+p3p+sweep)", "sweep32", "fused_pass", "shuffle", "Full", "Cost", "least_squares", "query",
         "LO minimal", "setup", "#Full", "#Cost"]
tot = np.zeros(12)
for q in range(Q):
    a = (int(off[q]) + 2000 - 32) & ~1
    tot += inl[a:a + 24].view(np.int64)
tot /= Q
print(f"Q={Q}: mean per query: lsq {res['lsq_calls'].mean():.1f} passes {res['passes'].mean():.1f} lo {res['num_lo'].mean():.2f}")
for n, v in zip(names, tot):
    if n.startswith("#"):
        print(f"  {n:28s} {v:10.1f}")
    else:
        print(f"  {n:28s} {v/1e3:10.1f} kcycles  {100*v/tot[7]:5.1f}%")
