"""API misuse / concurrency probe, run in a child process with faulthandler: every case must end in an error code or a
result, never in a crash. Usage: python tools/probes/misuse_probe.py"""
import subprocess
import sys

CASE = r'''
import os, sys, threading, ctypes as C
sys.path.insert(0, os.getcwd())
import numpy as np
import ransaclib_b200 as rlb
from ransaclib_b200 import synth, _capi
lib = _capi.load()
inst = synth.pose_instance(5, 400, 400)
f, p2, p3 = inst["focal"], inst["points2D"], inst["points3D"]
models = synth.perturbed_poses(3, inst["R_gt"], inst["c_gt"], 600)

def attempt(name, fn):
    try:
        r = fn()
        print("ok   ", name, type(r).__name__, flush=True)
    except Exception as e:
        print("error", name, type(e).__name__, str(e)[:90], flush=True)

ctx = rlb.Context(0)
sc = ctx.pose_scorer(p2, p3, f, f)
attempt("precision 99", lambda: sc.score_models(models, 144.0, precision=99))
attempt("thr 0", lambda: sc.score_models(models, 0.0))
attempt("thr negative", lambda: sc.score_models(models, -1.0))
attempt("thr nan", lambda: sc.score_models(models, float("nan")))
attempt("thr inf", lambda: sc.score_models(models, float("inf")))
for prec in (rlb.RLB_F32, rlb.RLB_F64, rlb.RLB_F64_FILTERED, rlb.RLB_F32_SORTED, rlb.RLB_F32_UNSORTED, rlb.RLB_F64_PLAIN):
    attempt("thr nan prec %d" % prec, lambda: sc.score_models(models, float("nan"), precision=prec))
    attempt("thr inf prec %d" % prec, lambda: sc.score_models(models, float("inf"), precision=prec))
    attempt("thr 0 prec %d" % prec, lambda: sc.score_models(models, 0.0, precision=prec))
    attempt("k=1 prec %d" % prec, lambda: sc.score_models(models[:1], 144.0, precision=prec))
    attempt("k=0 prec %d" % prec, lambda: sc.score_models(models[:0], 144.0, precision=prec))
    attempt("all nan models prec %d" % prec, lambda: sc.score_models(np.full((600, 12), np.nan), 144.0, precision=prec))
    attempt("huge models prec %d" % prec, lambda: sc.score_models(models * 1e200, 144.0, precision=prec))
attempt("inliers thr nan", lambda: sc.get_inliers(models[0], float("nan"), 144.0))
attempt("inliers thr inf", lambda: sc.get_inliers(models[0], float("inf"), float("inf")))
attempt("inliers nan model", lambda: sc.get_inliers(np.full(12, np.nan), 144.0, 144.0))
attempt("null model", lambda: _capi.check(lib.rlb_get_inliers(sc._h, None, None, None, None, None, None, None)))
attempt("null scorer", lambda: _capi.check(lib.rlb_get_inliers(None, None, None, None, None, None, None, None)))
e = ctx.pose_scorer(np.zeros((0, 2)), np.zeros((0, 3)), f, f)
for prec in (rlb.RLB_F32, rlb.RLB_F64, rlb.RLB_F64_FILTERED, rlb.RLB_F32_SORTED):
    attempt("empty data prec %d" % prec, lambda: e.score_models(models, 144.0, precision=prec))
attempt("empty data inliers", lambda: e.get_inliers(models[0], 144.0, 144.0))
attempt("samples out of range", lambda: sc.hypothesize_and_score(np.full((8, 4), 10**6, dtype=np.int32), 144.0))
attempt("samples negative", lambda: sc.hypothesize_and_score(np.full((8, 4), -3, dtype=np.int32), 144.0))
attempt("samples repeated", lambda: sc.hypothesize_and_score(np.zeros((8, 4), dtype=np.int32), 144.0))
ctx.close()
attempt("use after close", lambda: sc.score_models(models, 144.0))

# two threads, one context each, sweeping concurrently (the GIL is released inside the calls)
errs = []
def worker(seed):
    try:
        c = rlb.Context(0)
        s = c.pose_scorer(p2, p3, f, f)
        ref = None
        for it in range(40):
            for prec in (rlb.RLB_F64, rlb.RLB_F32, rlb.RLB_F64_FILTERED):
                r = s.score_models(models, 144.0, precision=prec)[0]
            g = s.get_inliers(models[seed], 144.0, 144.0)
            if ref is None:
                ref = r.copy()
            elif not np.array_equal(ref, r):
                errs.append("thread %d: results changed" % seed)
        s.close(); c.close()
    except Exception as ex:
        errs.append(repr(ex))
ts = [threading.Thread(target=worker, args=(i,)) for i in range(4)]
[t.start() for t in ts]; [t.join() for t in ts]
print("threads", "ok" if not errs else errs[:3], flush=True)
print("done", flush=True)
'''
r = subprocess.run([sys.executable, "-X", "faulthandler", "-c", CASE], capture_output=True, text=True, timeout=600)
print("rc", r.returncode)
print(r.stdout[-6000:])
print(r.stderr[-2500:])
