"""Tensor-level wrappers for the occupancy-grid update / compaction entry points of the C ABI
and the (single) collective of the hot path: a MAX all-reduce of `occs` across ranks."""
from __future__ import annotations

import ctypes as C
from typing import Tuple

import torch
from torch import Tensor

from . import _C
from ._grid_ops import _pinned_totals
from .pack import pack_info


def dist_info(group=None) -> Tuple[int, int]:
    """(world_size, rank) within `group` (None = the default process group); (1, 0) when
    torch.distributed is not initialised."""
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized():
        return dist.get_world_size(group), dist.get_rank(group)
    return 1, 0


def shard_range(n: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous slice [lo, hi) of n items owned by `rank`."""
    per = (n + world - 1) // world
    lo = min(rank * per, n)
    return lo, min(lo + per, n)


def all_reduce_max(occs: Tensor, group=None) -> None:
    """The only collective on the path: every rank decayed the same cells and
    max-ed in its own shard of evaluations; MAX across ranks completes the scatter-max."""
    import torch.distributed as dist

    dist.all_reduce(occs, op=dist.ReduceOp.MAX, group=group)


def cell_points(cells: Tensor, u: Tensor, resolution: Tensor, aabb: Tensor) -> Tensor:
    """World positions of jittered cell samples (occ_grid.py:385-392)."""
    n = cells.shape[0]
    # (a host list from the caller: `.tolist()` on a device tensor is a host round trip per call)
    res = [int(v) for v in (resolution.tolist() if isinstance(resolution, Tensor) else resolution)]
    dev = _C.require_cuda(cells, u, aabb)
    x = torch.empty((n, 3), dtype=torch.float32, device=dev)
    if n:
        cells, u, aabb = cells.long().contiguous(), u.float().contiguous(), aabb.float().contiguous()
        res_host = (C.c_int32 * 3)(*res)
        with torch.cuda.device(dev):
            rc = _C.lib().nfb_occ_cell_points(cells.data_ptr(), u.data_ptr(), n, res_host,
                                              aabb.data_ptr(), x.data_ptr(), _C.stream_ptr(dev))
        _C.check(rc, "occ_cell_points")
    return x


def select_cells(occs_lvl: Tensor, uniform: Tensor, binaries_lvl: Tensor) -> Tuple[Tensor, int, int]:
    """(cells, k1, k2): cells[:k1] = uniform[occs_lvl[uniform] >= 0], cells[k1:k1+k2] =
    nonzero(binaries_lvl.flatten()) — the two selections of occ_grid.py:344-365 in two launches and
    ONE host read-back (the reference: two boolean selects with a read-back each)."""
    dev = _C.require_cuda(occs_lvl, uniform, binaries_lvl)
    n, n_cells = uniform.shape[0], binaries_lvl.numel()
    assert occs_lvl.is_contiguous() and occs_lvl.dtype == torch.float32 and occs_lvl.numel() == n_cells
    uniform = uniform.long().contiguous()
    bins = binaries_lvl.contiguous()
    assert bins.dtype == torch.bool
    out = torch.empty((n + n_cells,), dtype=torch.long, device=dev)
    totals = _pinned_totals(dev)
    with torch.cuda.device(dev):
        lib = _C.lib()
        ws = _C.scratch(dev, int(lib.nfb_compact_workspace_bytes(max(n, n_cells))), "select")
        st = _C.stream_ptr(dev)
        t0, t1 = totals.data_ptr(), totals.data_ptr() + 8
        if n:  # (an empty tensor has no address to tell the two modes apart)
            _C.check(lib.nfb_occ_select_cells(occs_lvl.data_ptr(), uniform.data_ptr(), None, n, None,
                                              out.data_ptr(), t0, ws.data_ptr(), ws.numel(), st), "occ_select_cells")
        if n_cells:
            _C.check(lib.nfb_occ_select_cells(None, None, bins.data_ptr(), n_cells, t0 if n else None,
                                              out.data_ptr(), t1, ws.data_ptr(), ws.numel(), st), "occ_select_cells")
        torch.cuda.current_stream(dev).synchronize()
    return out, int(totals[0]) if n else 0, int(totals[1]) if n_cells else 0


def ema_update(occs: Tensor, cell_ids: Tensor, occ: Tensor, lo: int, hi: int, decay: float) -> None:
    """In place: occs[c] = max(occs[c] * decay, new values of c) for the cells in cell_ids;
    `occ` holds the evaluations of cell_ids[lo:hi]."""
    n = cell_ids.shape[0]
    if n == 0:
        return
    dev = _C.require_cuda(occs, cell_ids, occ)
    assert occs.is_contiguous() and occs.dtype == torch.float32
    cell_ids = cell_ids.long().contiguous()
    occ = occ.float().contiguous()
    assert occ.shape[0] == hi - lo, (occ.shape, lo, hi)
    tmp = torch.empty((n,), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        rc = _C.lib().nfb_occ_ema_update(occs.data_ptr(), cell_ids.data_ptr(), n, occ.data_ptr(), lo, hi,
                                         float(decay), tmp.data_ptr(), _C.stream_ptr(dev))
    _C.check(rc, "occ_ema_update")


def binarize(occs: Tensor, occ_thre: float) -> Tensor:
    """bool tensor occs > min(mean(occs[occs >= 0]), occ_thre) (occ_grid.py:403-404)."""
    dev = _C.require_cuda(occs)
    o = occs.contiguous()
    out = torch.empty(o.shape, dtype=torch.bool, device=dev)
    with torch.cuda.device(dev):
        need = int(_C.lib().nfb_occ_binarize_workspace_bytes())
        ws = _C.scratch(dev, need, "binarize")
        rc = _C.lib().nfb_occ_binarize(o.data_ptr(), o.numel(), float(occ_thre), out.data_ptr(), None,
                                       ws.data_ptr(), ws.numel(), _C.stream_ptr(dev))
    _C.check(rc, "occ_binarize")
    return out


def compact_samples(masks: Tensor, ray_indices: Tensor, t_starts: Tensor, t_ends: Tensor, n_rays: int):
    """(ray_indices[masks], t_starts[masks], t_ends[masks]) in one pass and one host sync
    (the reference: three boolean-mask selects, occ_grid.py:216-220)."""
    n = masks.numel()
    dev = _C.require_cuda(masks, ray_indices, t_starts, t_ends)
    if n == 0:
        return ray_indices, t_starts, t_ends
    o_ray, o_ts, o_te = torch.empty_like(ray_indices), torch.empty_like(t_starts), torch.empty_like(t_ends)
    totals = _pinned_totals(dev)
    with torch.cuda.device(dev):
        need = int(_C.lib().nfb_compact_workspace_bytes(n))
        ws = _C.scratch(dev, need, "compact")
        rc = _C.lib().nfb_compact_samples(masks.data_ptr(), ray_indices.data_ptr(), t_starts.data_ptr(),
                                          t_ends.data_ptr(), n, o_ray.data_ptr(), o_ts.data_ptr(),
                                          o_te.data_ptr(), totals.data_ptr(), ws.data_ptr(), ws.numel(),
                                          _C.stream_ptr(dev))
        _C.check(rc, "compact_samples")
        torch.cuda.current_stream(dev).synchronize()
    k = int(totals[0])
    return o_ray[:k], o_ts[:k], o_te[:k]


def visibility_compact(from_density: bool, values: Tensor, t_starts: Tensor, t_ends: Tensor,
                       packed_info: Tensor, ray_indices: Tensor, early_stop_eps: float, alpha_thre: float,
                       want_mask: bool = False):
    """render_visibility_from_{density,alpha} + the three boolean selects of occ_grid.py:186-220
    + pack_info of the survivors, as one kernel and one host sync.  `values` = sigmas (density
    mode) or alphas; `packed_info` must be canonical and dense over the samples.
    Returns (ray_indices', t_starts', t_ends', packed_info', mask or None)."""
    n = values.numel()
    n_rays = packed_info.shape[0]
    dev = _C.require_cuda(values, t_starts, t_ends, packed_info, ray_indices)
    if n == 0:
        return ray_indices, t_starts, t_ends, packed_info, (torch.zeros(0, dtype=torch.bool, device=dev) if want_mask else None)
    f32 = lambda t: t.contiguous() if t.dtype == torch.float32 else t.float().contiguous()
    values, t_starts, t_ends = f32(values), f32(t_starts), f32(t_ends)
    ray_indices = ray_indices.contiguous()
    assert ray_indices.dtype == torch.long and ray_indices.shape == values.shape == t_starts.shape == t_ends.shape
    o_ray, o_ts, o_te = torch.empty_like(ray_indices), torch.empty_like(t_starts), torch.empty_like(t_ends)
    o_pi = torch.empty_like(packed_info)
    mask = torch.empty(values.shape, dtype=torch.bool, device=dev) if want_mask else None
    totals = _pinned_totals(dev)
    with torch.cuda.device(dev):
        ws, ws_bytes, epoch = _C.workspace(dev, n)
        rc = _C.lib().nfb_visibility_compact(
            int(from_density), None if from_density else values.data_ptr(), t_starts.data_ptr(),
            t_ends.data_ptr(), values.data_ptr() if from_density else None, None, packed_info.data_ptr(),
            n_rays, n, ray_indices.data_ptr(), float(early_stop_eps), float(alpha_thre), o_ray.data_ptr(),
            o_ts.data_ptr(), o_te.data_ptr(), o_pi.data_ptr(), None if mask is None else mask.data_ptr(),
            totals.data_ptr(), ws.data_ptr(), ws_bytes, epoch, _C.stream_ptr(dev))
        _C.check(rc, "visibility_compact")
        torch.cuda.current_stream(dev).synchronize()
    k = int(totals[0])
    return o_ray[:k], o_ts[:k], o_te[:k], o_pi, mask


def mark_invisible_cells(occs: Tensor, cams: Tensor, resolution: Tensor, levels: int, aabbs: Tensor,
                         width: float, height: float, near_plane: float) -> None:
    """In place (occ_grid.py:263-332): occs = 0 for visible cells, -1 for the rest; cells already
    at -1 are left alone.  cams: [n_cams, 21] = (w2c_R | w2c_T | K) per camera."""
    dev = _C.require_cuda(occs, cams, aabbs)
    assert occs.is_contiguous() and occs.dtype == torch.float32 and cams.shape[-1] == 21
    res_host = (C.c_int32 * 3)(*[int(v) for v in (resolution.tolist() if isinstance(resolution, Tensor) else resolution)])
    aabbs = aabbs.float().contiguous()
    with torch.cuda.device(dev):
        rc = _C.lib().nfb_mark_invisible_cells(cams.data_ptr(), cams.shape[0], res_host, int(levels),
                                               aabbs.data_ptr(), float(width), float(height), float(near_plane),
                                               occs.data_ptr(), _C.stream_ptr(dev))
    _C.check(rc, "mark_invisible_cells")
