"""The reference's own Python package running on its own CUDA kernels (test infrastructure, GPU only).

`baseline/_ref/nerfacc` is the unmodified reference installed with
    BUILD_NO_CUDA=1 pip install --no-index --no-build-isolation --no-deps --target baseline/_ref <copy of /root/reference>
(git-ignored, travels with gpurun); its compiled backend is oracle/_ref/nerfacc_cuda.so (built
from the reference sources by oracle/build_ref.py), injected as `nerfacc.csrc`, the name the
reference's loader tries first (nerfacc/cuda/_backend.py:52-54) — so nothing is JIT-compiled.
Used where a check needs the reference's host logic too (PropNetEstimator, OccGridEstimator).
"""
import os
import sys

import torch

from tests import _refcuda

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "baseline", "_ref")
_mod = None


def available() -> bool:
    return os.path.isdir(os.path.join(PKG, "nerfacc")) and _refcuda.available()


def load():
    """Returns the reference package (module `nerfacc` as the reference wrote it)."""
    global _mod
    if _mod is None:
        assert available()
        csrc = _refcuda.mod()
        prev = sys.modules.get("nerfacc")
        assert prev is None or getattr(prev, "__file__", "").startswith(PKG), "another `nerfacc` is already imported"
        sys.modules["nerfacc.csrc"] = csrc
        sys.path.insert(0, PKG)
        try:
            import nerfacc

            nerfacc.csrc = csrc
            from nerfacc.cuda import _backend  # noqa: F401  (binds _C = nerfacc.csrc now)

            assert _backend._C is csrc
        finally:
            sys.path.remove(PKG)
        _mod = nerfacc
    return _mod
