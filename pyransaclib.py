"""Top-level ``pyransaclib`` module, as the reference installs it (examples/pyransaclib.cc:38-46 built by
setup.py:55-110): ``import pyransaclib; pyransaclib.ransaclib_localization(...)`` works unchanged with this
repository on ``sys.path`` (or with this one file copied next to the ``ransaclib_b200`` package). The
implementation is ``ransaclib_b200.pyransaclib`` (CUDA through the C ABI; no CPU fallback)."""
from ransaclib_b200.pyransaclib import ransaclib_localization, ransaclib_localization_batch  # noqa: F401

__all__ = ["ransaclib_localization", "ransaclib_localization_batch"]
