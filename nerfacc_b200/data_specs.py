"""RaySamples / RayIntervals — the packed-ray data model (reference nerfacc/data_specs.py:12-180).

Public fields are identical to the reference's dataclasses.  The reference marshals them to a
C++ `RaySegmentsSpec` (stack/unbind of packed_info at every crossing); here the kernels take
`packed_info` as stored, so there is nothing to marshal.
"""
from dataclasses import dataclass
from typing import Optional

import torch


@dataclass
class RaySamples:
    """Ray samples, batched `vals` (n_rays, n_samples) or flattened `vals` (all_samples,).

    Flattened data needs `packed_info` (n_rays, 2) = (start, count) per ray and/or
    `ray_indices` (all_samples,).  `is_valid` marks the filled slots of an over-allocated
    layout (traverse_grids with over_allocate=True).
    """

    vals: torch.Tensor
    packed_info: Optional[torch.Tensor] = None
    ray_indices: Optional[torch.Tensor] = None
    is_valid: Optional[torch.Tensor] = None

    @property
    def device(self) -> torch.device:
        return self.vals.device

    @property
    def is_batched(self) -> bool:
        return self.vals.dim() > 1


@dataclass
class RayIntervals:
    """Ray intervals as a list of edges, batched `vals` (n_rays, n_edges) or flattened.

    For flattened data `is_left[i]` / `is_right[i]` tell whether edge i is the left / right
    edge of some interval, which represents both continuous and non-continuous intervals:
    interval k of a ray spans [vals[is_left][k], vals[is_right][k]].
    """

    vals: torch.Tensor
    packed_info: Optional[torch.Tensor] = None
    ray_indices: Optional[torch.Tensor] = None
    is_left: Optional[torch.Tensor] = None
    is_right: Optional[torch.Tensor] = None

    @property
    def device(self) -> torch.device:
        return self.vals.device

    @property
    def is_batched(self) -> bool:
        return self.vals.dim() > 1
