"""Object lifetimes across the C ABI: a scorer points into its context (stream, memory pool), and a garbage
collector finalises the two in no particular order. rlb_ctx_destroy on a context with live scorers defers the
destruction to the last rlb_scorer_destroy; these tests pin that (a script with module-level functions that left a
context and a scorer alive used to die with SIGSEGV at interpreter exit)."""
import subprocess
import sys

import numpy as np
import pytest

from ransaclib_b200 import synth

pytestmark = pytest.mark.gpu

THR = 144.0


def test_context_destroyed_before_its_scorer():
    import ransaclib_b200 as rlb

    inst = synth.pose_instance(11, 300, 300)
    f = inst["focal"]
    ctx = rlb.Context(0)
    sc = ctx.pose_scorer(inst["points2D"], inst["points3D"], f, f)
    models = synth.perturbed_poses(3, inst["R_gt"], inst["c_gt"], 16)
    want, _, _, _ = sc.score_models(models, THR)
    # the context first, straight through the C ABI (what an unordered finaliser does)
    assert ctx._lib.rlb_ctx_destroy(ctx._h) == 0
    ctx._h = None
    got, _, _, _ = sc.score_models(models, THR)  # the context lives on while the scorer does
    np.testing.assert_array_equal(got, want)
    sc.close()  # completes the deferred destruction
    # and the device is still usable afterwards
    ctx2 = rlb.Context(0)
    sc2 = ctx2.pose_scorer(inst["points2D"], inst["points3D"], f, f)
    again, _, _, _ = sc2.score_models(models, THR)
    np.testing.assert_array_equal(again, want)
    sc2.close()
    ctx2.close()


SCRIPT = r"""
import os, sys
sys.path.insert(0, os.getcwd())
import ransaclib_b200 as rlb
from ransaclib_b200 import synth

inst = synth.pose_instance(12, 300, 300)
f = inst["focal"]
ctx = rlb.Context(0)
sc = ctx.pose_scorer(inst["points2D"], inst["points3D"], f, f)
models = synth.perturbed_poses(3, inst["R_gt"], inst["c_gt"], 16)


def score():  # a module-level function: globals <-> function cycle, so ctx and sc go through the cyclic collector
    return sc.score_models(models, 144.0)[0]


print("scores", float(score().min()), flush=True)
# neither closed: interpreter exit finalises them in whatever order the collector picks
"""


def test_interpreter_exit_with_live_objects():
    r = subprocess.run([sys.executable, "-X", "faulthandler", "-c", SCRIPT], capture_output=True, text=True,
                       timeout=300)
    assert r.returncode == 0, (r.returncode, r.stdout[-500:], r.stderr[-2000:])
    assert "scores" in r.stdout
