"""Timeline of one batched-localization launch (debug_flags bit 1): when each query began / ended (global ns
timer), on which SM / CTA, to see how the makespan splits into steady state and tail.
usage: timeline_localize.py Q [mode]"""
import os
import sys

import numpy as np

sys.path.insert(0, ".")
if len(sys.argv) > 2:
    os.environ["RLB_LOC_MODE"] = sys.argv[2]
import ransaclib_b200 as rlb
from ransaclib_b200 import synth

Q = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
base = 64
p2s, p3s, fs = [], [], []
for q in range(base):
    inst = synth.pose_instance(5000 + q, 1000, 1000)
    p2s.append(inst["points2D"]); p3s.append(inst["points3D"]); fs.append(float(np.float32(inst["focal"])))
reps = (Q + base - 1) // base
p2 = np.concatenate(p2s * reps)[: 2000 * Q]
p3 = np.concatenate(p3s * reps)[: 2000 * Q]
f = np.array((fs * reps)[:Q])
off = np.arange(Q + 1, dtype=np.int64) * 2000
seeds = np.arange(42, 42 + Q, dtype=np.uint32)
ctx = rlb.Context(0)
o = ctx.localize_options(12.0, 10, 100, 10000)
o.debug_flags = 2
ctx.localize_batch(off, p2, p3, f, f, seeds, o)
poses, res, inl = ctx.localize_batch(off, p2, p3, f, f, seeds, o)
rec = np.zeros((Q, 16), np.int64)
for q in range(Q):
    a = (int(off[q]) + 2000 - 32) & ~1
    rec[q] = inl[a:a + 32].view(np.int64)
t0 = rec[:, 12].min()
beg, end = (rec[:, 12] - t0) * 1e-6, (rec[:, 13] - t0) * 1e-6  # ms
dur = end - beg
print(f"Q={Q}: makespan {end.max():.2f} ms; query duration mean {dur.mean():.2f} median {np.median(dur):.2f} "
      f"p90 {np.percentile(dur, 90):.2f} max {dur.max():.2f} ms; sum/592 = {dur.sum()/592:.2f} ms")
print("num_lo histogram:", np.bincount(res["num_lo"]).tolist())
for lo in range(1, res["num_lo"].max() + 1):
    m = res["num_lo"] == lo
    if m.any():
        print(f"  lo={lo}: n={m.sum():5d} mean duration {dur[m].mean():6.2f} ms  lsq {res['lsq_calls'][m].mean():6.1f}")
# active queries over time
ts = np.linspace(0, end.max(), 21)
act = [(int(((beg <= t) & (end > t)).sum())) for t in ts]
print("active queries at 5% steps of the makespan:", act)
late = np.argsort(end)[-8:]
print("last finishers: ", [(int(q), f"{beg[q]:.1f}->{end[q]:.1f}", int(res['num_lo'][q])) for q in late])
sm_busy = {}
for q in range(Q):
    sm_busy.setdefault(int(rec[q, 14]), []).append(dur[q])
tot = np.array([sum(v) for v in sm_busy.values()])
print(f"SMs used {len(sm_busy)}; per-SM summed query time min {tot.min():.1f} mean {tot.mean():.1f} max {tot.max():.1f} ms")
