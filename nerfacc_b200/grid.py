"""Grid traversal — drop-in for nerfacc/grid.py (reference nerfacc/grid.py:14-237)."""
from typing import Optional, Tuple

import torch
from torch import Tensor

from . import _C
from ._grid_ops import Marcher
from ._packing import mark_canonical, remember_packed_info
from .data_specs import RayIntervals, RaySamples


@torch.no_grad()
def ray_aabb_intersect(
    rays_o: Tensor,
    rays_d: Tensor,
    aabbs: Tensor,
    near_plane: float = -float("inf"),
    far_plane: float = float("inf"),
    miss_value: float = float("inf"),
) -> Tuple[Tensor, Tensor, Tensor]:
    """Ray-AABB intersection (slab test) of every ray with every box.

    Args:
        rays_o: (n_rays, 3) ray origins.  rays_d: (n_rays, 3) normalized directions.
        aabbs: (m, 6) boxes {xmin, ymin, zmin, xmax, ymax, zmax}.
        near_plane / far_plane: clip range for hits.  miss_value: t for misses.

    Returns:
        t_mins (n_rays, m), t_maxs (n_rays, m), hits (n_rays, m) bool.
    Reference: nerfacc/grid.py:14-51, csrc/grid.cu:284-313, 477-519.
    """
    assert rays_o.ndim == 2 and rays_o.shape[-1] == 3
    assert rays_d.ndim == 2 and rays_d.shape[-1] == 3
    assert aabbs.ndim == 2 and aabbs.shape[-1] == 6
    dev = _C.require_cuda(rays_o, rays_d, aabbs)
    o, d, bb = rays_o.float().contiguous(), rays_d.float().contiguous(), aabbs.float().contiguous()
    n, m = o.shape[0], bb.shape[0]
    t_mins = torch.empty((n, m), dtype=torch.float32, device=dev)
    t_maxs = torch.empty((n, m), dtype=torch.float32, device=dev)
    hits = torch.empty((n, m), dtype=torch.bool, device=dev)
    if n * m > 0:
        with torch.cuda.device(dev):
            rc = _C.lib().nfb_ray_aabb_intersect(
                o.data_ptr(), d.data_ptr(), bb.data_ptr(), n, m, float(near_plane), float(far_plane),
                float(miss_value), t_mins.data_ptr(), t_maxs.data_ptr(), hits.data_ptr(),
                _C.stream_ptr(dev))
        _C.check(rc, "ray_aabb_intersect")
    return t_mins, t_maxs, hits


def _ray_aabb_intersect(
    rays_o: Tensor,
    rays_d: Tensor,
    aabbs: Tensor,
    near_plane: float = -float("inf"),
    far_plane: float = float("inf"),
    miss_value: float = float("inf"),
) -> Tuple[Tensor, Tensor, Tensor]:
    """Pure-PyTorch slab test with the semantics of :func:`ray_aabb_intersect` (any device).
    Reference: nerfacc/grid.py:54-90."""
    lo, hi = aabbs[None, :, :3], aabbs[None, :, 3:]
    o, d = rays_o[:, None, :], rays_d[:, None, :]
    ta, tb = (lo - o) / d, (hi - o) / d
    t_mins = torch.minimum(ta, tb).amax(dim=-1)
    t_maxs = torch.maximum(ta, tb).amin(dim=-1)
    hits = (t_maxs > t_mins) & (t_maxs > 0)
    t_mins = torch.where(hits, t_mins.clamp(near_plane, far_plane), miss_value)
    t_maxs = torch.where(hits, t_maxs.clamp(near_plane, far_plane), miss_value)
    return t_mins, t_maxs, hits


@torch.no_grad()
def traverse_grids(
    # rays
    rays_o: Tensor,  # [n_rays, 3]
    rays_d: Tensor,  # [n_rays, 3]
    # grids
    binaries: Tensor,  # [m, resx, resy, resz]
    aabbs: Tensor,  # [m, 6]
    # options
    near_planes: Optional[Tensor] = None,  # [n_rays]
    far_planes: Optional[Tensor] = None,  # [n_rays]
    step_size: Optional[float] = 1e-3,
    cone_angle: Optional[float] = 0.0,
    traverse_steps_limit: Optional[int] = None,
    over_allocate: Optional[bool] = False,
    rays_mask: Optional[Tensor] = None,  # [n_rays]
    # pre-compute intersections
    t_sorted: Optional[Tensor] = None,  # [n_rays, n_grids * 2]
    t_indices: Optional[Tensor] = None,  # [n_rays, n_grids * 2]
    hits: Optional[Tensor] = None,  # [n_rays, n_grids]
) -> Tuple[RayIntervals, RaySamples, Tensor]:
    """Ray traversal within multiple (nested or not) occupancy grids.  Not differentiable.

    Args:
        rays_o, rays_d: (n_rays, 3) origins and normalized directions.
        binaries: (m, resx, resy, resz) bool grids of one resolution; aabbs: (m, 6).
        near_planes / far_planes: (n_rays,) optional, default 0 / infinity.
        step_size: marching step (<= 0: one interval per occupied voxel crossing).
        cone_angle: > 0 grows the step linearly with t.
        traverse_steps_limit: max samples per ray (required with over_allocate).
        over_allocate: single pass into `limit`-strided slots; compact with is_left /
            is_right / is_valid; `rays_mask` (n_rays,) skips rays in this mode.
        t_sorted, t_indices, hits: optional precomputed sorted ray/box events.

    Returns:
        (RayIntervals, RaySamples, termination_planes (n_rays,)).
    Reference: nerfacc/grid.py:94-192, csrc/grid.cu:320-474.
    """
    limit = -1 if traverse_steps_limit is None else int(traverse_steps_limit)
    if over_allocate:
        assert limit > 0, "traverse_steps_limit must be set if over_allocate is True."
    m = Marcher(rays_o, rays_d, binaries, aabbs,
                0.0 if near_planes is None else near_planes,
                float("inf") if far_planes is None else far_planes,
                step_size, cone_angle, limit, rays_mask=rays_mask if over_allocate else None,
                t_sorted=t_sorted, t_indices=t_indices, hits=hits)
    dev, n_rays = m.dev, m.n_rays
    if not over_allocate:
        sample_info, interval_info, term, n_samples, n_edges = m.count(True, True)
        iv, sm = m.fill_full(False, sample_info, interval_info, n_samples, n_edges)
    else:
        alive = torch.ones((n_rays,), dtype=torch.long, device=dev) if rays_mask is None else rays_mask.long()
        alloc = alive * limit
        sample_starts = torch.cumsum(alloc, 0) - alloc
        interval_starts = sample_starts * 2
        n_alloc = int(alive.sum().item()) * limit
        sample_info = torch.empty((n_rays, 2), dtype=torch.long, device=dev)
        interval_info = torch.empty((n_rays, 2), dtype=torch.long, device=dev)
        term = torch.empty((n_rays,), dtype=torch.float32, device=dev)
        iv, sm = m.fill_full(True, sample_info, interval_info, n_alloc, 2 * n_alloc,
                             sample_starts, interval_starts, term)
    mark_canonical(sample_info, total=None if over_allocate else n_samples)
    mark_canonical(interval_info, total=None if over_allocate else n_edges)
    if not over_allocate:
        remember_packed_info(sm["ray_indices"], sample_info, n_rays)
    intervals = RayIntervals(vals=iv["vals"], packed_info=interval_info, ray_indices=iv["ray_indices"],
                             is_left=iv["is_left"], is_right=iv["is_right"])
    samples = RaySamples(vals=sm["vals"], packed_info=sample_info, ray_indices=sm["ray_indices"],
                         is_valid=sm["is_valid"])
    return intervals, samples, term


def _enlarge_aabb(aabb, factor: float) -> Tensor:
    """Scale a box about its centre.  Reference: nerfacc/grid.py:195-198."""
    half = (aabb[3:] - aabb[:3]) / 2
    mid = (aabb[:3] + aabb[3:]) / 2
    return torch.cat([mid - half * factor, mid + half * factor])


def _query(x: Tensor, data: Tensor, base_aabb: Tensor) -> Tuple[Tensor, Tensor]:
    """Look up nested grids (each level 2x the previous about the same centre) at points x.

    Returns (values at the finest level containing each point, mask of points inside any level).
    Test oracle of the reference (nerfacc/grid.py:201-237), restated.
    """
    lo, hi = base_aabb[:3], base_aabb[3:]
    centred = (x - lo) / (hi - lo) - 0.5  # base box -> [-0.5, 0.5]^3
    # frexp(|c|_inf) picks the level; a floor of 0.1 keeps frexp away from 0 (exponent 0 there)
    reach = centred.abs().amax(dim=-1).clamp_min(0.1)
    mip = (torch.frexp(reach)[1].long() + 1).clamp_min(0)
    inside = mip < data.shape[0]
    unit = centred / (2**mip)[:, None] + 0.5
    res = torch.tensor(data.shape[1:], device=x.device)
    ijk = torch.minimum((unit * res).long(), res - 1)
    mip = mip.clamp_max(data.shape[0] - 1)
    return data[mip, ijk[:, 0], ijk[:, 1], ijk[:, 2]] * inside, inside
