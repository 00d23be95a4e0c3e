"""Book-keeping for packed ray layouts.

Two facts about tensors this package itself produced are remembered (weakly, keyed by tensor
identity and in-place version counter) so later calls can skip work without changing results:

* a ``packed_info`` is *canonical* (starts == exclusive cumsum of counts, as produced by
  ``pack_info`` / ``traverse_grids`` / ``OccGridEstimator.sampling``): the flat look-back
  kernels may be used instead of the general one-warp-per-ray kernels;
* the ``packed_info`` that belongs to a ``ray_indices`` tensor: ``rendering`` and friends take
  ``ray_indices`` (the public API) but the kernels want the 16 B/ray form, and rebuilding it
  costs a pass over 8 B/sample (reference: nerfacc/volrend.py:200-201, 256-257).
"""
from __future__ import annotations

import weakref
from typing import Dict, Optional, Tuple

import torch

_CANONICAL: Dict[int, Tuple[weakref.ref, int, Optional[int]]] = {}
_HEAD_BITS: Dict[int, Tuple[weakref.ref, int, torch.Tensor]] = {}
_PACKED_OF: Dict[int, Tuple[weakref.ref, int, torch.Tensor, int]] = {}


def _drop(table, key):
    def cb(_ref):
        table.pop(key, None)
    return cb


def mark_canonical(packed_info: torch.Tensor, total: Optional[int] = None) -> torch.Tensor:
    """`total`: set when the layout is also dense — the chunks tile [0, total) without gaps
    (not the case for over-allocated traversals)."""
    key = id(packed_info)
    _CANONICAL[key] = (weakref.ref(packed_info, _drop(_CANONICAL, key)), packed_info._version,
                       None if total is None else int(total))
    return packed_info


def dense_total(packed_info: torch.Tensor) -> Optional[int]:
    ent = _CANONICAL.get(id(packed_info))
    if ent is None or ent[0]() is not packed_info or ent[1] != packed_info._version:
        return None
    return ent[2]


def head_bits(packed_info: torch.Tensor, n: int) -> Optional[torch.Tensor]:
    """The 1-bit-per-sample ray boundaries of a dense canonical `packed_info` over n samples
    (nfb_head_bits), built on first use and kept as long as the packed_info tensor lives;
    None when the layout is not known to be dense."""
    if dense_total(packed_info) != int(n) or n == 0:
        return None
    key = id(packed_info)
    ent = _HEAD_BITS.get(key)
    if ent is not None and ent[0]() is packed_info and ent[1] == packed_info._version:
        return ent[2]
    from . import _C

    dev = packed_info.device
    lib = _C.lib()
    bits = torch.empty((int(lib.nfb_head_bits_bytes(n)) + 3) // 4, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        rc = lib.nfb_head_bits(packed_info.data_ptr(), packed_info.shape[0], n, bits.data_ptr(),
                               _C.stream_ptr(dev))
    _C.check(rc, "head_bits")
    _HEAD_BITS[key] = (weakref.ref(packed_info, _drop(_HEAD_BITS, key)), packed_info._version, bits)
    return bits


def is_canonical(packed_info: torch.Tensor) -> bool:
    ent = _CANONICAL.get(id(packed_info))
    return ent is not None and ent[0]() is packed_info and ent[1] == packed_info._version


def remember_packed_info(ray_indices: torch.Tensor, packed_info: torch.Tensor, n_rays: int) -> None:
    key = id(ray_indices)
    _PACKED_OF[key] = (weakref.ref(ray_indices, _drop(_PACKED_OF, key)), ray_indices._version,
                       packed_info, int(n_rays))


def recall_packed_info(ray_indices: torch.Tensor, n_rays: Optional[int]) -> Optional[torch.Tensor]:
    ent = _PACKED_OF.get(id(ray_indices))
    if ent is None or ent[0]() is not ray_indices or ent[1] != ray_indices._version:
        return None
    if n_rays is not None and int(n_rays) != ent[3]:
        return None
    return ent[2]


def as_packed_info(packed_info: torch.Tensor) -> Tuple[torch.Tensor, bool]:
    """Returns (int64 contiguous [n_rays, 2] tensor, canonical?)."""
    assert (
        packed_info.dim() == 2 and packed_info.shape[-1] == 2
    ), "packed_info must be 2-D with shape (B, 2)."
    canon = is_canonical(packed_info)
    if packed_info.dtype != torch.long or not packed_info.is_contiguous():
        packed_info = packed_info.long().contiguous()
    return packed_info, canon
