"""Smallest case of every kernel family, for compute-sanitizer (memcheck / racecheck / synccheck):
  python tools/sanitize_cases.py [tc|ffma2|f64|line|inlier|solvers|localize|localize_warp|line_lomsac|all]
No torch: numpy + the C ABI only (host-pointer entry points), so the tool sees just this library's kernels."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ransaclib_b200 as rlb
from ransaclib_b200 import synth
from ransaclib_b200.sampling import UniformSampling

which = sys.argv[1] if len(sys.argv) > 1 else "all"
ctx = rlb.Context(0)


def want(name):
    return which in ("all", name)


inst = synth.pose_instance(3, 700, 500)
p2, p3, f = inst["points2D"], inst["points3D"], inst["focal"]
if want("tc") or want("ffma2") or want("f64") or want("inlier") or want("solvers"):
    sc = ctx.pose_scorer(p2, p3, f, f)
    if want("tc"):  # K >= 512: tensor-core sweep (tcgen05 / TMEM / TMA ring)
        s = UniformSampling(1, len(p2), 4).sample_fast(640)
        r = sc.hypothesize_and_score(s, 144.0, precision=rlb.RLB_F32)
        print("tc sweep: plan", ctx.last_sweep_plan(), "best", r["best_id"], flush=True)
    if want("ffma2"):  # K < 512: FFMA2 sweep (TMA bulk ring)
        s = UniformSampling(2, len(p2), 4).sample_fast(200)
        r = sc.hypothesize_and_score(s, 144.0, precision=rlb.RLB_F32)
        print("ffma2 sweep: plan", ctx.last_sweep_plan(), "best", r["best_id"], flush=True)
    if want("f64"):
        s = UniformSampling(3, len(p2), 4).sample_fast(300)
        r = sc.hypothesize_and_score(s, 144.0, precision=rlb.RLB_F64)
        print("f64 sweep: plan", ctx.last_sweep_plan(), "best", r["best_id"], flush=True)
    if want("inlier") or want("solvers"):
        s = UniformSampling(4, len(p2), 4).sample_fast(64)
        r = sc.hypothesize_and_score(s, 144.0, precision=rlb.RLB_F64)
        lists, counts, score = sc.get_inliers(r["best_model"], 144.0, 144.0)
        print("inlier pass:", counts, score, flush=True)
    sc.close()
if want("line"):
    xy = synth.line_instance(5, 1500, 1500)
    sc = ctx.line_scorer(xy)
    s = UniformSampling(5, len(xy), 2).sample_fast(700)
    r = sc.hypothesize_and_score(s, 1e-4, precision=rlb.RLB_F32)
    r = sc.hypothesize_and_score(s[:90], 1e-4, precision=rlb.RLB_F64)
    lists, counts, score = sc.get_inliers(r["best_model"], 1e-4, 1e-4)
    print("line sweeps + inlier pass:", counts, flush=True)
    sc.close()
if want("localize") or want("localize_warp"):
    if which == "localize_warp":
        os.environ["RLB_LOC_MODE"] = "0"
    qs = [synth.pose_instance(70 + q, 150 + 40 * q, 150) for q in range(6)]
    off = np.concatenate([[0], np.cumsum([len(q["points2D"]) for q in qs])]).astype(np.int64)
    P2 = np.concatenate([q["points2D"] for q in qs])
    P3 = np.concatenate([q["points3D"] for q in qs])
    F = np.array([float(np.float32(q["focal"])) for q in qs])
    opt = ctx.localize_options(12.0, 10, 100, 10000)
    poses, res, inl = ctx.localize_batch(off, P2, P3, F, F, np.arange(3, 9, dtype=np.uint32), opt)
    print("localize:", res["num_inliers"].tolist(), res["num_iterations"].tolist(), flush=True)
if want("line_lomsac"):
    xs = [synth.line_instance(90 + q, 300, 200 + 100 * q) for q in range(3)]
    loff = np.concatenate([[0], np.cumsum([len(x) for x in xs])]).astype(np.int64)
    lines, lres, linl = ctx.line_lomsac_batch(loff, np.concatenate(xs), np.arange(11, 14, dtype=np.uint32),
                                              ctx.lomsac_options(sq_thr=1e-4))
    print("line_lomsac:", lres["num_inliers"].tolist(), flush=True)
ctx.close()
print("done", which)
