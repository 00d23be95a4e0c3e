"""The C-ABI library loads without a GPU and exports every symbol include/nerfacc_b200.h declares;
argument errors are reported through return codes + nfb_last_error (no compute calls here)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    with open(os.path.join(ROOT, "include", "nerfacc_b200.h")) as f:
        src = f.read()
    return sorted(set(re.findall(r"NFB_API\s+[\w\s\*]+?\b(nfb_\w+)\s*\(", src)))


def test_header_and_loader_agree():
    from nerfacc_b200 import _C

    hdr = _header_symbols()
    assert len(hdr) >= 20
    assert hdr == _C.declared_symbols(), (set(hdr) ^ set(_C.declared_symbols()))


def test_library_exports_every_declared_symbol():
    from nerfacc_b200 import _C, _build

    _C.lib()
    raw = ctypes.CDLL(_build.LIB)
    for name in _header_symbols():
        assert hasattr(raw, name), f"{name} missing from libnerfacc_b200.so"
    assert _C.lib().nfb_compiled_arch() == 100
    assert _C.lib().nfb_version() >= 100
    assert _C.lib().nfb_workspace_bytes(1 << 24) > 0


def test_argument_errors_use_return_codes():
    from nerfacc_b200 import _C

    l = _C.lib()
    rc = l.nfb_scan_f32(None, None, None, 4, 16, 7, 0, 0, 0, 1, None, 0, 1, None)  # unknown op
    assert rc == 1 and "op" in _C.last_error()
    rc = l.nfb_accumulate_fwd(None, None, 0, None, None, 1, 1, None, 0, None, 0, 1, None)  # dim 0
    assert rc == 1
    assert l.nfb_scan_f32(None, None, None, 0, 0, 0, 0, 0, 0, 1, None, 0, 1, None) == 0  # empty: no-op


def test_no_cpu_fallback():
    import torch

    import nerfacc_b200 as nf

    x = torch.rand(9)
    pi = torch.tensor([[0, 2], [2, 3], [5, 4]])
    with pytest.raises(RuntimeError, match="CUDA"):
        nf.inclusive_sum(x, pi)
    with pytest.raises(NotImplementedError):
        nf.pack_info(torch.tensor([0, 0, 1]))
    with pytest.raises(RuntimeError, match="CUDA"):
        nf.ray_aabb_intersect(torch.rand(4, 3), torch.rand(4, 3), torch.rand(2, 6))


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "nerfacc_b200")
    for dp, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h")):
                with open(os.path.join(dp, fn)) as f:
                    src = f.read()
                assert not re.search(r"^\s*(import|from)\s+oracle\b", src, re.M), fn
                assert "nfb_oracle" not in src, fn
