"""CPU tests of the boundary: the C-ABI library loads, exports every symbol that
include/ransaclib_b200.h declares, and fails loudly (no CPU fallback) without a GPU."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "ransaclib_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rlb_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from ransaclib_b200 import _capi

    lib = _capi.load()
    declared = _declared()
    assert len(declared) >= 25
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/ransaclib_b200.h but not exported"
    assert sorted(_capi.SYMBOLS) == declared
    assert lib.rlb_version() >= 100


def test_no_cpu_fallback():
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import ransaclib_b200 as rlb

    with pytest.raises(rlb.RlbError, match="no CPU fallback"):
        rlb.Context(0)


def test_product_does_not_touch_the_oracle():
    """Nothing under ransaclib_b200/ or include/ may reference oracle/."""
    bad = []
    for base in ("ransaclib_b200", "include"):
        for dp, _, fs in os.walk(os.path.join(ROOT, base)):
            for f in fs:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".cc", ".cpp")):
                    txt = open(os.path.join(dp, f), errors="ignore").read()
                    if re.search(r"oracle[/_]|liboracle|libransac_ref", txt):
                        bad.append(os.path.join(dp, f))
    assert not bad, bad


def test_top_level_pyransaclib_module_mirrors_the_reference():
    """`import pyransaclib` (examples/pyransaclib.cc:38-46): same function name and keyword arguments."""
    import inspect

    import pyransaclib

    sig = inspect.signature(pyransaclib.ransaclib_localization)
    want = ["query_name", "focal_x", "focal_y", "points2D_vec", "points3D_vec", "inlier_threshold",
            "number_lo_steps", "min_num_iterations", "max_num_iterations"]
    assert list(sig.parameters)[:len(want)] == want
    assert callable(pyransaclib.ransaclib_localization_batch)
