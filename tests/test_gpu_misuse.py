"""Bad arguments, degenerate inputs and concurrent contexts must end in an error code or a result, never in a crash:
tools/probes/misuse_probe.py runs the cases in a child process (non-finite and zero thresholds in every precision
mode, empty data sets, k = 0 / 1, non-finite and huge models, out-of-range samples, use after close, four threads
with a context each sweeping concurrently)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_misuse_never_crashes():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "probes", "misuse_probe.py")], cwd=ROOT,
                       capture_output=True, text=True, timeout=900)
    out = r.stdout
    assert r.returncode == 0 and "\nrc 0" in "\n" + out, out[-3000:] + r.stderr[-2000:]
    assert "threads ok" in out and "\ndone" in out
    for must_fail in ("precision 99", "null model", "null scorer", "samples out of range", "samples negative",
                      "use after close"):
        assert "error " + must_fail in out, must_fail
    for must_work in ("thr inf prec 0", "thr 0 prec 1", "k=1 prec 3", "all nan models prec 4", "empty data inliers",
                      "samples repeated"):
        assert "ok    " + must_work in out, must_work
