import os, sys, time
sys.path.insert(0, os.getcwd())
import numpy as np
import ransaclib_b200 as rlb
from ransaclib_b200 import synth
hp = synth.hybrid_pose_instance(33, 25000, 25000, 10000, 10000)
rows0 = np.ascontiguousarray(np.concatenate([hp["points2D"], hp["points3D"]], axis=1))
rows1 = np.ascontiguousarray(np.concatenate([hp["cam1"], hp["world1"]], axis=1))
src = """__device__ double rlb_eval_t(const double* m, int type, const double* p) {
  const double* X = (type == 0) ? p + 2 : p + 3;
  const double d0 = X[0] - m[9], d1 = X[1] - m[10], d2 = X[2] - m[11];
  const double pz = m[2] * d0 + m[5] * d1 + m[8] * d2;
  const double px = m[0] * d0 + m[3] * d1 + m[6] * d2;
  const double py = m[1] * d0 + m[4] * d1 + m[7] * d2;
  if (type == 0) { if (pz < 0.0) return 1e308; const double du = px / pz * 185.0 - p[0], dv = py / pz * 185.0 - p[1]; return du*du+dv*dv; }
  const double ex = px - p[0], ey = py - p[1], ez = pz - p[2];
  return ex*ex+ey*ey+ez*ez;
}"""
ctx = rlb.Context(0)
def t(fn, reps=20):
    fn(); t0 = time.perf_counter()
    for _ in range(reps): fn()
    return (time.perf_counter() - t0) / reps * 1e6
def mk():
    s = ctx.custom_hybrid_scorer(rows0, rows1, 12, src); s.close()
print("custom hybrid create+close us", round(t(mk), 1))
def mk2():
    s = ctx.pose_scorer(hp["points2D"], hp["points3D"], hp["focal"], hp["focal"]); s.close()
print("pose create+close us", round(t(mk2), 1))
s = ctx.custom_hybrid_scorer(rows0, rows1, 12, src)
m = synth.pose_to_model(hp["R_gt"], hp["c_gt"])
print("custom pass with list us", round(t(lambda: s.get_inliers(m, [144.0, 0.0025], [144.0, 0.0025]), 100), 1))
models = synth.perturbed_poses(3, hp["R_gt"], hp["c_gt"], 256)
print("custom score 256 us", round(t(lambda: s.score_models(models, [144.0, 0.0025]), 50), 1))
s.close(); ctx.close()

# throughput of a user-defined functor on a large sweep: the pose residual as source, K = 16384 x N = 1M, f64
import torch
inst = synth.pose_instance(9000, 500000, 500000)
rows = np.ascontiguousarray(np.concatenate([inst["points2D"], inst["points3D"]], axis=1))
src1 = src.replace("rlb_eval_t(const double* m, int type, const double* p)", "rlb_eval(const double* m, const double* p)").replace("(type == 0) ? p + 2 : p + 3", "p + 2").replace("if (type == 0) {", "{")
ctx = rlb.Context(0)
ctx.set_timing(True)
models = synth.perturbed_poses(3, inst["R_gt"], inst["c_gt"], 16384)
try:
    cs = ctx.custom_scorer(rows, 12, src1)
    cs.score_models(models, 144.0)
    ms = []
    for _ in range(3):
        cs.score_models(models, 144.0)
        ms.append(ctx.kernel_time_ms(0))
    print("custom functor sweep 16384 x 1M f64: %.2f ms -> %.3e evals/s" % (min(ms), 16384 * 1e6 / (min(ms) * 1e-3)))
    src2 = src1.replace("px / pz * 185.0", "px * inv * 185.0").replace("py / pz * 185.0", "py * inv * 185.0").replace(
        "const double du =", "const double inv = rlb_rcp(pz); const double du =")
    cr = ctx.custom_scorer(rows, 12, src2)
    cr.score_models(models, 144.0)
    ms = []
    for _ in range(3):
        cr.score_models(models, 144.0)
        ms.append(ctx.kernel_time_ms(0))
    print("the same with rlb_rcp instead of the two divisions: %.2f ms -> %.3e evals/s" % (min(ms), 16384 * 1e6 / (min(ms) * 1e-3)))
    ps = ctx.pose_scorer(inst["points2D"], inst["points3D"], inst["focal"], inst["focal"])
    ps.score_models(models, 144.0, precision=rlb.RLB_F64_PLAIN)
    ms = []
    for _ in range(3):
        ps.score_models(models, 144.0, precision=rlb.RLB_F64_PLAIN)
        ms.append(ctx.kernel_time_ms(0))
    print("built-in plain f64 sweep, same size: %.2f ms -> %.3e evals/s" % (min(ms), 16384 * 1e6 / (min(ms) * 1e-3)))
except Exception as e:
    print("large sweep probe failed:", repr(e)[:300])
