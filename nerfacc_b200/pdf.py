"""Inverse-CDF resampling and per-ray searchsorted — drop-in for nerfacc/pdf.py
(reference nerfacc/pdf.py:13-139: same names, arguments, return types and id conventions).

Both ops go through the C ABI of libnerfacc_b200 (`nfb_importance_sampling`, `nfb_searchsorted`,
csrc/pdf.cu): one warp per ray, the ray's rows staged in shared memory, centres and edges written
by one kernel.  Stratified sampling draws its one-offset-per-ray from torch's CUDA generator here
(`1 - torch.rand`, the (0,1] range of curand_uniform in csrc/pdf.cu:139-145) and hands it to the
kernel as a tensor, so a fixed seed reproduces and tests can pin the offsets.
"""
from typing import Optional, Tuple, Union

import torch
from torch import Tensor

from . import _C
from ._packing import as_packed_info, mark_canonical
from .data_specs import RayIntervals, RaySamples

Segments = Union[RayIntervals, RaySamples]


def _rows(seg: Segments, what: str):
    """-> (contiguous fp32 vals, packed_info or None, n_rays, row length (batched) or 0)."""
    vals = seg.vals
    if vals.dtype != torch.float32:
        raise RuntimeError(f"{what}.vals must be float32, got {vals.dtype}")
    vals = vals.contiguous()
    if vals.dim() > 1:
        n = vals.shape[-1]
        return vals, None, (vals.numel() // n if n > 0 else 0), n
    if seg.packed_info is None:
        raise ValueError(f"{what}: flattened vals need packed_info")
    info, _ = as_packed_info(seg.packed_info)
    return vals, info, info.shape[0], 0


def searchsorted(sorted_sequence: Segments, values: Segments) -> Tuple[Tensor, Tensor]:
    """Per-ray searchsorted on batched or flattened tensors (reference nerfacc/pdf.py:13-63).

    Returns LongTensors `ids_left`, `ids_right` shaped like `values.vals` with
    `sorted_sequence.vals[ids_left] <= values.vals < sorted_sequence.vals[ids_right]`, queries
    outside a ray's range treated as clipped to it.  Batched `values`: ids index the last dim of
    `sorted_sequence.vals`; flattened `values`: ids index the flattened `sorted_sequence.vals`.

    >>> seq = RayIntervals(vals=torch.tensor([0.0, 1.0, 0.0, 1.0, 2.0], device="cuda"),
    ...                    packed_info=torch.tensor([[0, 2], [2, 3]], device="cuda"))
    >>> q = RayIntervals(vals=torch.tensor([0.5, 1.5, 2.5], device="cuda"),
    ...                  packed_info=torch.tensor([[0, 1], [1, 2]], device="cuda"))
    >>> searchsorted(seq, q)
    (tensor([0, 3, 3], device='cuda:0'), tensor([1, 4, 4], device='cuda:0'))
    """
    q_vals, q_info, n_rays_q, n_q = _rows(values, "values")
    k_vals, k_info, n_rays_k, n_k = _rows(sorted_sequence, "sorted_sequence")
    dev = _C.require_cuda(q_vals, k_vals, q_info, k_info)
    ids_left = torch.empty(q_vals.shape, dtype=torch.long, device=dev)
    ids_right = torch.empty(q_vals.shape, dtype=torch.long, device=dev)
    if q_vals.numel() == 0:
        return ids_left, ids_right
    if n_rays_q != n_rays_k:
        raise ValueError(f"values cover {n_rays_q} rays but sorted_sequence {n_rays_k}")
    with torch.cuda.device(dev):
        rc = _C.lib().nfb_searchsorted(
            q_vals.data_ptr(), _C.ptr(q_info), n_q, k_vals.data_ptr(), _C.ptr(k_info), n_k, n_k,
            n_rays_q, ids_left.data_ptr(), ids_right.data_ptr(), _C.stream_ptr(dev))
    _C.check(rc, "searchsorted")
    return ids_left, ids_right


def _stratified_bias(n_rays: int, device: torch.device) -> Tensor:
    # one offset per ray in (0, 1], csrc/pdf.cu:139-145
    return 1.0 - torch.rand(n_rays, device=device, dtype=torch.float32)


def importance_sampling(
    intervals: RayIntervals,
    cdfs: Tensor,
    n_intervals_per_ray: Union[Tensor, int],
    stratified: bool = False,
    *,
    _bias: Optional[Tensor] = None,
    _samples: bool = True,
    _stot: Optional[Tuple[str, float, float]] = None,
) -> Tuple[RayIntervals, RaySamples]:
    """Inverse-transform resampling of every ray (reference nerfacc/pdf.py:66-139).

    `intervals` holds the edges along each ray (batched or flattened) and `cdfs` the CDF at those
    edges.  Each ray gets `n_intervals_per_ray` new samples at the centres of equal-mass bins
    (one common random offset per ray when `stratified`), and the matching intervals.

    int `n_intervals_per_ray`: `intervals.vals` (n_rays, n+1), `samples.vals` (n_rays, n).
    Tensor (n_rays,): flattened outputs with `packed_info`, `ray_indices` (and `is_left`,
    `is_right` on the intervals); rays asking for 0 samples get no edges.

    >>> intervals = RayIntervals(vals=torch.tensor([0.0, 1.0, 0.0, 1.0, 2.0], device="cuda"),
    ...                          packed_info=torch.tensor([[0, 2], [2, 3]], device="cuda"))
    >>> cdfs = torch.tensor([0.0, 0.5, 0.0, 0.5, 1.0], device="cuda")
    >>> intervals, samples = importance_sampling(intervals, cdfs, 2)
    >>> intervals.vals
    tensor([[0.0000, 0.5000, 1.0000],
            [0.0000, 1.0000, 2.0000]], device='cuda:0')
    >>> samples.vals
    tensor([[0.2500, 0.7500],
            [0.5000, 1.5000]], device='cuda:0')
    """
    vals, in_info, n_rays, n_in = _rows(intervals, "intervals")
    if cdfs.dtype != torch.float32:
        raise RuntimeError(f"cdfs must be float32, got {cdfs.dtype}")
    cdfs = cdfs.contiguous()
    assert cdfs.numel() == vals.numel(), "cdfs must have the same number of elements as intervals.vals"
    per_ray = isinstance(n_intervals_per_ray, Tensor)
    counts = n_intervals_per_ray.contiguous().long().reshape(-1) if per_ray else None
    dev = _C.require_cuda(vals, cdfs, in_info, counts)
    if per_ray and counts.numel() != n_rays:
        raise ValueError(f"n_intervals_per_ray has {counts.numel()} entries for {n_rays} rays")
    bias = None
    if _bias is not None:
        bias = _bias.to(device=dev, dtype=torch.float32).contiguous().reshape(-1)
        assert bias.numel() == n_rays
    elif stratified:
        bias = _stratified_bias(n_rays, dev)

    if not per_ray:
        n_out = int(n_intervals_per_ray)
        lead = vals.shape[:-1] if vals.dim() > 1 else (n_rays,)
        # callers that only want the intervals (PropNetEstimator) skip the centres' write-out
        want_samples = _samples or vals.dim() == 1
        centres = torch.empty(lead + (n_out,), dtype=torch.float32, device=dev) if want_samples else None
        edges = torch.empty(lead + (n_out + 1,), dtype=torch.float32, device=dev)
        if _stot is not None:
            # PropNetEstimator: the edges in s and, from the same kernel, mapped to ray distance
            # (third return value); None when this shape has no fused path
            t_edges = None
            kind, t_min, t_max = _stot
            fused = (vals.dim() > 1 and in_info is None and not want_samples and kind in ("uniform", "lindisp")
                     and isinstance(t_min, (int, float)) and isinstance(t_max, (int, float)) and n_in <= 1024)
            if fused and n_rays > 0 and n_out > 0:
                c0, c1 = (float(t_min), float(t_max)) if kind == "uniform" else (1 / t_min, 1 / t_max)
                t_edges = torch.empty_like(edges)
                with torch.cuda.device(dev):
                    rc = _C.lib().nfb_importance_sampling_stot(
                        vals.data_ptr(), cdfs.data_ptr(), n_rays, n_in, n_out, _C.ptr(bias), edges.data_ptr(),
                        t_edges.data_ptr(), 1 if kind == "uniform" else 2, c0, c1, _C.stream_ptr(dev))
                _C.check(rc, "importance_sampling_stot")
                return RayIntervals(vals=edges), RaySamples(vals=None), t_edges
        if n_rays > 0 and n_out > 0:
            with torch.cuda.device(dev):
                rc = _C.lib().nfb_importance_sampling(
                    vals.data_ptr(), cdfs.data_ptr(), _C.ptr(in_info), n_rays, n_in, n_in,
                    None, n_out, n_out, None, None, _C.ptr(bias),
                    edges.data_ptr(), _C.ptr(centres), _C.stream_ptr(dev))
            _C.check(rc, "importance_sampling")
        if _stot is not None:
            return RayIntervals(vals=edges), RaySamples(vals=centres), None
        return RayIntervals(vals=edges), RaySamples(vals=centres)

    # per-ray counts: flattened outputs.  The totals size the allocations (one host sync, as the
    # reference's memalloc_data_from_chunk does, data_spec.hpp:86-96).
    e_counts = (counts + 1) * (counts > 0)
    both = torch.stack([counts, e_counts], -1)
    ends = torch.cumsum(both, 0)
    s_info = mark_canonical(torch.stack([ends[:, 0] - counts, counts], -1).contiguous())
    e_info = mark_canonical(torch.stack([ends[:, 1] - e_counts, e_counts], -1).contiguous())
    n_s, n_e = (int(v) for v in ends[-1].tolist()) if n_rays > 0 else (0, 0)
    centres = torch.empty(n_s, dtype=torch.float32, device=dev)
    edges = torch.empty(n_e, dtype=torch.float32, device=dev)
    ray_ids = torch.arange(n_rays, device=dev)
    s_rays = torch.repeat_interleave(ray_ids, counts, output_size=n_s)
    e_rays = torch.repeat_interleave(ray_ids, e_counts, output_size=n_e)
    is_left = torch.ones(n_e, dtype=torch.bool, device=dev)
    is_right = torch.ones(n_e, dtype=torch.bool, device=dev)
    has = e_counts > 0
    is_right[e_info[has, 0]] = False
    is_left[e_info[has, 0] + e_counts[has] - 1] = False
    if n_s > 0:
        with torch.cuda.device(dev):
            rc = _C.lib().nfb_importance_sampling(
                vals.data_ptr(), cdfs.data_ptr(), _C.ptr(in_info), n_rays, n_in, n_in,
                counts.data_ptr(), 0, 0, e_info.data_ptr(), s_info.data_ptr(), _C.ptr(bias),
                edges.data_ptr(), centres.data_ptr(), _C.stream_ptr(dev))
        _C.check(rc, "importance_sampling")
    return (
        RayIntervals(vals=edges, packed_info=e_info, ray_indices=e_rays, is_left=is_left, is_right=is_right),
        RaySamples(vals=centres, packed_info=s_info, ray_indices=s_rays),
    )


def _sample_from_weighted(
    bins: Tensor,
    weights: Tensor,
    num_samples: int,
    stratified: bool = False,
    vmin: float = -torch.inf,
    vmax: float = torch.inf,
) -> Tuple[Tensor, Tensor]:
    """Plain-torch resampler kept for API parity (reference nerfacc/pdf.py:142-224; its tests use
    it as the slow cross-check of importance_sampling).  bins (..., B+1), weights (..., B) ->
    edges (..., S+1), centres (..., S)."""
    n_bins, s = weights.shape[-1], num_samples
    assert bins.shape[-1] == n_bins + 1
    eps = torch.finfo(weights.dtype).eps
    pdf = torch.nn.functional.normalize(weights, p=1, dim=-1)
    zero, one = torch.zeros_like(pdf[..., :1]), torch.ones_like(pdf[..., :1])
    cdf = torch.cat([zero, torch.cumsum(pdf[..., :-1], dim=-1), one], dim=-1)
    kw = dict(dtype=bins.dtype, device=bins.device)
    if stratified:
        u_max = eps + (1 - eps) / s
        jitter = torch.rand(*bins.shape[:-1], 1, **kw) * ((1 - u_max) / (s - 1) - eps)
        u = torch.linspace(0, 1 - u_max, s, **kw) + jitter
    else:
        pad = 1 / (2 * s)
        u = torch.linspace(pad, 1 - pad - eps, s, **kw).broadcast_to(bins.shape[:-1] + (s,))
    hi = torch.searchsorted(cdf.contiguous(), u.contiguous(), side="right")
    idx = torch.cat([hi - 1, hi], dim=-1)
    c0, c1 = cdf.gather(-1, idx).split(s, dim=-1)
    b0, b1 = bins.gather(-1, idx).split(s, dim=-1)
    centres = b0 + (u - c0) / torch.clamp(c1 - c0, min=eps) * (b1 - b0)
    mids = (centres[..., 1:] + centres[..., :-1]) / 2
    first = (2 * centres[..., :1] - mids[..., :1]).clamp_min(vmin)
    last = (2 * centres[..., -1:] - mids[..., -1:]).clamp_max(vmax)
    return torch.cat([first, mids, last], dim=-1), centres
