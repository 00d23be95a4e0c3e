"""Thin tensor-level wrappers over the fused rendering entry points of the C ABI
(nfb_render_weight_fwd/bwd, nfb_accumulate_fwd/bwd) and the autograd Functions built on them.

The public, reference-compatible API lives in volrend.py; this module only marshals
tensors to pointers.  Segment layout is passed as either a canonical int64 packed_info
or `uniform_len` (batched [..., n] tensors flattened, every row one segment).
"""
from __future__ import annotations

import os
from typing import Optional, Tuple

import torch
from torch import Tensor

from . import _C
from ._packing import head_bits


def _p(t: Optional[Tensor]):
    return None if t is None else t.data_ptr()


def weight_fwd(from_density: bool, alphas: Optional[Tensor], t_starts: Optional[Tensor],
               t_ends: Optional[Tensor], sigmas: Optional[Tensor], prefix_trans: Optional[Tensor],
               packed_info: Optional[Tensor], uniform_len: int, want_weights: bool,
               want_trans: bool, want_alphas: bool, vis: Optional[Tuple[float, float]] = None):
    """Returns (weights, trans, alphas_out, vis_mask); entries not asked for are None."""
    like = sigmas if from_density else alphas
    dev = _C.require_cuda(like, packed_info)
    n = like.numel()
    mk = lambda want: torch.empty_like(like) if want else None
    weights, trans, alphas_out = mk(want_weights), mk(want_trans), mk(want_alphas and from_density)
    mask = torch.empty(like.shape, dtype=torch.bool, device=dev) if vis is not None else None
    if n > 0:
        eps, thre = (float(vis[0]), float(vis[1])) if vis is not None else (0.0, 0.0)
        n_rays = packed_info.shape[0] if packed_info is not None else n // max(uniform_len, 1)
        bits = head_bits(packed_info, n) if packed_info is not None else None
        with torch.cuda.device(dev):
            ws, ws_bytes, epoch = _C.workspace(dev, n)
            if bits is not None:  # dense layout: warp tiles
                rc = _C.lib().nfb_render_weight_bits_fwd(
                    int(from_density), _p(alphas), _p(t_starts), _p(t_ends), _p(sigmas),
                    _p(prefix_trans), _p(bits), n, _p(weights), _p(trans), _p(alphas_out), _p(mask),
                    eps, thre, ws.data_ptr(), ws_bytes, epoch, _C.stream_ptr(dev))
            else:
                rc = _C.lib().nfb_render_weight_fwd(
                    int(from_density), _p(alphas), _p(t_starts), _p(t_ends), _p(sigmas),
                    _p(prefix_trans), _p(packed_info), n_rays, int(uniform_len), n, _p(weights),
                    _p(trans), _p(alphas_out), _p(mask), eps, thre, ws.data_ptr(), ws_bytes, epoch,
                    _C.stream_ptr(dev))
        _C.check(rc, "render_weight_fwd")
    return weights, trans, alphas_out, mask


def weight_bwd(from_density: bool, g_weights: Optional[Tensor], g_trans: Optional[Tensor],
               g_alphas: Optional[Tensor], alphas: Optional[Tensor], t_starts: Optional[Tensor],
               t_ends: Optional[Tensor], sigmas: Optional[Tensor], trans: Tensor,
               packed_info: Optional[Tensor], uniform_len: int) -> Tensor:
    like = sigmas if from_density else alphas
    dev = _C.require_cuda(like, trans)
    n = like.numel()
    g_in = torch.empty_like(like)
    if n > 0:
        n_rays = packed_info.shape[0] if packed_info is not None else n // max(uniform_len, 1)
        bits = head_bits(packed_info, n) if packed_info is not None else None
        with torch.cuda.device(dev):
            ws, ws_bytes, epoch = _C.workspace(dev, n)
            if bits is not None:
                rc = _C.lib().nfb_render_weight_bits_bwd(
                    int(from_density), _p(g_weights), _p(g_trans), _p(g_alphas), _p(alphas),
                    _p(t_starts), _p(t_ends), _p(sigmas), _p(trans), _p(bits), n, _p(g_in),
                    ws.data_ptr(), ws_bytes, epoch, _C.stream_ptr(dev))
            else:
                rc = _C.lib().nfb_render_weight_bwd(
                    int(from_density), _p(g_weights), _p(g_trans), _p(g_alphas), _p(alphas),
                    _p(t_starts), _p(t_ends), _p(sigmas), _p(trans), _p(packed_info), n_rays,
                    int(uniform_len), n, _p(g_in), ws.data_ptr(), ws_bytes, epoch, _C.stream_ptr(dev))
        _C.check(rc, "render_weight_bwd")
    return g_in


def _cg(g: Optional[Tensor]) -> Optional[Tensor]:
    return None if g is None else g.contiguous()


class WeightFromAlpha(torch.autograd.Function):
    """(weights, trans) = f(alphas): fused replacement of volrend.py:161-206, 267-309."""

    @staticmethod
    def forward(ctx, alphas, prefix_trans, packed_info, uniform_len: int, want_weights: bool):
        alphas = alphas.contiguous()
        weights, trans, _, _ = weight_fwd(False, alphas, None, None, None, prefix_trans,
                                          packed_info, uniform_len, want_weights, True, False)
        ctx.set_materialize_grads(False)
        ctx.uniform_len = uniform_len
        ctx.has_pi = packed_info is not None
        if ctx.needs_input_grad[0]:
            ctx.save_for_backward(alphas, trans, *([packed_info] if packed_info is not None else []))
        if weights is None:
            weights = trans.new_empty(0)
            ctx.mark_non_differentiable(weights)
        return weights, trans

    @staticmethod
    def backward(ctx, g_weights, g_trans):
        saved = ctx.saved_tensors
        alphas, trans = saved[0], saved[1]
        packed_info = saved[2] if ctx.has_pi else None
        if g_weights is None and g_trans is None:
            return torch.zeros_like(alphas), None, None, None, None
        g = weight_bwd(False, _cg(g_weights), _cg(g_trans), None, alphas, None, None, None,
                       trans, packed_info, ctx.uniform_len)
        return g, None, None, None, None


class WeightFromDensity(torch.autograd.Function):
    """(weights, trans, alphas) = f(sigmas): fused replacement of volrend.py:209-264, 312-362."""

    @staticmethod
    def forward(ctx, sigmas, t_starts, t_ends, prefix_trans, packed_info, uniform_len: int,
                want_weights: bool):
        sigmas, t_starts, t_ends = sigmas.contiguous(), t_starts.contiguous(), t_ends.contiguous()
        weights, trans, alphas, _ = weight_fwd(True, None, t_starts, t_ends, sigmas, prefix_trans,
                                               packed_info, uniform_len, want_weights, True, True)
        ctx.set_materialize_grads(False)
        ctx.uniform_len = uniform_len
        ctx.has_pi = packed_info is not None
        if ctx.needs_input_grad[0]:
            ctx.save_for_backward(sigmas, t_starts, t_ends, trans,
                                  *([packed_info] if packed_info is not None else []))
        if weights is None:
            weights = trans.new_empty(0)
            ctx.mark_non_differentiable(weights)
        return weights, trans, alphas

    @staticmethod
    def backward(ctx, g_weights, g_trans, g_alphas):
        saved = ctx.saved_tensors
        sigmas, t_starts, t_ends, trans = saved[:4]
        packed_info = saved[4] if ctx.has_pi else None
        if g_weights is None and g_trans is None and g_alphas is None:
            return (torch.zeros_like(sigmas),) + (None,) * 6
        g = weight_bwd(True, _cg(g_weights), _cg(g_trans), _cg(g_alphas), None, t_starts, t_ends,
                       sigmas, trans, packed_info, ctx.uniform_len)
        return (g,) + (None,) * 6


# ----------------------------------------------------------------------------- accumulate
def accumulate_fwd(weights: Tensor, values: Optional[Tensor], packed_info: Optional[Tensor],
                   ray_indices: Optional[Tensor], n_rays: int, out: Optional[Tensor] = None) -> Tensor:
    dev = _C.require_cuda(weights, values, packed_info, ray_indices, out)
    dim = 1 if values is None else values.shape[-1]
    add_into = out is not None
    if out is None:
        out = torch.empty((n_rays, dim), dtype=weights.dtype, device=dev)
    n = weights.numel()
    bits = None
    if (n > 0 and 2 <= dim <= 4 and packed_info is not None and ray_indices is not None
            and _per_ray_kernels(n, n_rays)):
        # dense layout: warp tiles (csrc/render_bits.cu); measured at 36.4M samples: D = 3 0.22 ms vs
        # 0.33 ms on the flat CTA tiles, D = 1 0.106 vs 0.098 ms (kept on the flat tiles)
        bits = head_bits(packed_info, n)
    if bits is not None and n_rays > 0:
        with torch.cuda.device(dev):
            ws, ws_bytes, epoch = _C.workspace(dev, 8 * n)
            rc = _C.lib().nfb_accumulate_bits_fwd(
                _p(weights), _p(values), dim, _p(ray_indices), _p(bits), n_rays, n, out.data_ptr(),
                int(add_into), ws.data_ptr(), ws_bytes, epoch, _C.stream_ptr(dev))
        _C.check(rc, "accumulate_bits_fwd")
    elif n_rays > 0 and dim > 0:
        with torch.cuda.device(dev):
            ws, ws_bytes, epoch = _C.workspace(dev, max(n, 1))
            rc = _C.lib().nfb_accumulate_fwd(
                _p(weights), _p(values), dim, _p(packed_info), _p(ray_indices), n_rays, n,
                out.data_ptr(), int(add_into), ws.data_ptr(), ws_bytes, epoch, _C.stream_ptr(dev))
        _C.check(rc, "accumulate_fwd")
    return out


def accumulate_bwd(g_out: Tensor, weights: Tensor, values: Optional[Tensor], ray_indices: Tensor,
                   need_gw: bool, need_gv: bool):
    dev = _C.require_cuda(g_out, weights, values, ray_indices)
    dim = g_out.shape[-1]
    g_w = torch.empty_like(weights) if need_gw else None
    g_v = torch.empty_like(values) if (need_gv and values is not None) else None
    n = weights.numel()
    if n > 0 and (g_w is not None or g_v is not None):
        with torch.cuda.device(dev):
            rc = _C.lib().nfb_accumulate_bwd(g_out.data_ptr(), _p(weights), _p(values), dim,
                                             ray_indices.data_ptr(), n, _p(g_w), _p(g_v),
                                             _C.stream_ptr(dev))
        _C.check(rc, "accumulate_bwd")
    return g_w, g_v


class Accumulate(torch.autograd.Function):
    """out[n_rays, D] = segmented sum of weights[:, None] * values (volrend.py:532-547)."""

    @staticmethod
    def forward(ctx, weights, values, packed_info, ray_indices, n_rays: int):
        weights = weights.contiguous()
        values = None if values is None else values.contiguous()
        out = accumulate_fwd(weights, values, packed_info, ray_indices, n_rays)
        ctx.save_for_backward(weights, *([values] if values is not None else []), ray_indices)
        ctx.has_values = values is not None
        return out

    @staticmethod
    def backward(ctx, g_out):
        saved = ctx.saved_tensors
        weights = saved[0]
        values = saved[1] if ctx.has_values else None
        ray_indices = saved[-1]
        g_w, g_v = accumulate_bwd(g_out.contiguous(), weights, values, ray_indices,
                                  ctx.needs_input_grad[0], ctx.has_values and ctx.needs_input_grad[1])
        return g_w, g_v, None, None, None


# ----------------------------------------------------------------------------- rendering()
# Mean samples per ray up to which rendering() uses the per-ray kernels (a group of lanes walks
# one ray); above it the flat tile kernels, whose cost does not depend on how the samples are
# distributed over the rays.  NFB_RAYS_L forces the lanes per ray (4/8/16/32) for experiments.
_PER_RAY_MAX_MEAN = 512
_RAYS_L = int(os.environ.get("NFB_RAYS_L", "0"))


def _per_ray_kernels(n: int, n_rays: int) -> bool:
    return n_rays > 0 and n <= _PER_RAY_MAX_MEAN * n_rays


# NFB_RENDER = bits | rays | flat forces one of the three implementations (experiments, tests).
_RENDER_MODE = os.environ.get("NFB_RENDER", "")


def _render_mode(n: int, n_rays: int, packed_info: Tensor) -> str:
    """bits: dense canonical layout -> warp tiles (csrc/render_bits.cu), whatever the ray
    lengths; rays: a group of lanes per ray (any chunk layout); flat: CTA tiles with look-back
    for very long rays in a layout that is not known to be dense."""
    mode = _RENDER_MODE
    if mode not in ("bits", "rays", "flat"):
        mode = "bits"
    if mode == "bits" and _RENDER_MODE != "bits" and not _per_ray_kernels(n, n_rays):
        # rays of > 512 samples on average: a streaming tile that continues a ray waits for the
        # tile before it, a chain as long as the ray; the flat tiles' look-back has no such chain
        mode = "flat"
    if mode == "bits" and (n == 0 or n_rays == 0 or head_bits(packed_info, n) is None):
        mode = "rays" if _RENDER_MODE != "flat" else "flat"
    if mode == "rays" and _RENDER_MODE != "rays" and not _per_ray_kernels(n, n_rays):
        mode = "flat"
    return mode


class RenderComposite(torch.autograd.Function):
    """rendering() of volrend.py:14-158 for flattened CUDA inputs, one forward and one backward
    kernel.  Which pair depends on what is known about the layout (`_render_mode`): the streaming
    warp-tile kernels for dense canonical layouts (nfb_render_bits_fwd / _bwd), a group of lanes
    per ray for arbitrary chunks (nfb_render_rays_fwd / _bwd), or, for rays of more than 512
    samples on average, the flat CTA tiles — weights forward (nfb_render_weight_fwd), composite
    forward (nfb_composite_fwd), one backward (nfb_composite_bwd).

    Returns (colors_sum [R,3], opacity [R,1], depth [R,1], weights, trans, alphas).  With
    `norm_depth` depth is sum(w t_mid) / max(opacity, eps) (volrend.py:147-150), formed and
    differentiated by one launch over the rays each way (nfb_depth_normalize_fwd / _bwd); the
    background blend stays a torch op on [R, .] tensors in volrend.py.
    """

    @staticmethod
    def forward(ctx, field, rgbs, t_starts, t_ends, packed_info, ray_indices, n_rays: int,
                from_density: bool, norm_depth: bool = False):
        field, rgbs = field.contiguous(), rgbs.contiguous()
        t_starts, t_ends = t_starts.contiguous(), t_ends.contiguous()
        dev = _C.require_cuda(field, rgbs, t_starts, t_ends, packed_info, ray_indices)
        n = field.numel()
        # one [5, n_rays]-sized buffer: the bits kernel clears it with a single memset
        out5 = torch.empty(5 * n_rays, dtype=torch.float32, device=dev)
        colors = out5[: 3 * n_rays].view(n_rays, 3)
        opacity = out5[3 * n_rays: 4 * n_rays].view(n_rays, 1)
        depth = out5[4 * n_rays:].view(n_rays, 1)
        mode = _render_mode(n, n_rays, packed_info)
        per_ray = mode == "rays"
        if mode == "bits":
            # warp tiles over the dense layout (csrc/render_bits.cu)
            weights, trans = torch.empty_like(field), torch.empty_like(field)
            alphas = torch.empty_like(field) if from_density else None
            bits = head_bits(packed_info, n)
            with torch.cuda.device(dev):
                ws, ws_bytes, epoch = _C.workspace(dev, 8 * n)
                rc = _C.lib().nfb_render_bits_fwd(
                    int(from_density), None if from_density else _p(field), _p(t_starts), _p(t_ends),
                    _p(field) if from_density else None, _p(rgbs), _p(ray_indices), _p(bits), n_rays, n,
                    _p(weights), _p(trans), _p(alphas), colors.data_ptr(), opacity.data_ptr(),
                    depth.data_ptr(), ws.data_ptr(), ws_bytes, epoch, _C.stream_ptr(dev))
            _C.check(rc, "render_bits_fwd")
        elif per_ray:
            # one kernel: weights + the three per-ray sums (csrc/render_rays.cu)
            weights, trans = torch.empty_like(field), torch.empty_like(field)
            alphas = torch.empty_like(field) if from_density else None
            if n_rays > 0:
                with torch.cuda.device(dev):
                    rc = _C.lib().nfb_render_rays_fwd(
                        int(from_density), None if from_density else _p(field), _p(t_starts), _p(t_ends),
                        _p(field) if from_density else None, _p(rgbs), _p(packed_info), n_rays, n,
                        _p(weights), _p(trans), _p(alphas), colors.data_ptr(), opacity.data_ptr(),
                        depth.data_ptr(), _RAYS_L, _C.stream_ptr(dev))
                _C.check(rc, "render_rays_fwd")
        else:
            if from_density:
                weights, trans, alphas, _ = weight_fwd(True, None, t_starts, t_ends, field, None,
                                                       packed_info, 0, True, True, True)
            else:
                weights, trans, _, _ = weight_fwd(False, field, None, None, None, None, packed_info, 0,
                                                  True, True, False)
                alphas = None
            if n_rays > 0:
                with torch.cuda.device(dev):
                    ws, ws_bytes, epoch = _C.workspace(dev, max(n, 1))
                    rc = _C.lib().nfb_composite_fwd(
                        _p(weights), _p(rgbs), _p(t_starts), _p(t_ends), _p(packed_info), n_rays, n,
                        colors.data_ptr(), opacity.data_ptr(), depth.data_ptr(), ws.data_ptr(), ws_bytes,
                        epoch, _C.stream_ptr(dev))
                _C.check(rc, "composite_fwd")
        ctx.per_ray = per_ray
        ctx.mode = mode
        # the head bits travel with the node: the identity-keyed cache misses when saved-tensor
        # hooks (save_on_cpu, checkpoint wrappers) hand the backward a different packed_info object
        ctx.bits = bits if mode == "bits" else None
        ctx.set_materialize_grads(False)
        ctx.from_density = from_density
        ctx.n_rays = n_rays
        if norm_depth and n_rays > 0:
            # depth = sum(w t_mid) / max(opacity, eps) (volrend.py:147-150): one launch over the rays
            with torch.cuda.device(dev):
                rc = _C.lib().nfb_depth_normalize_fwd(opacity.data_ptr(), depth.data_ptr(), n_rays,
                                                      _C.stream_ptr(dev))
            _C.check(rc, "depth_normalize_fwd")
        ctx.norm_depth = bool(norm_depth)
        ctx.save_for_backward(field, rgbs, t_starts, t_ends, trans, packed_info, ray_indices,
                              *([opacity, depth] if norm_depth else []))
        if alphas is None:
            alphas = field.new_empty(0)
            ctx.mark_non_differentiable(alphas)
        return colors, opacity, depth, weights, trans, alphas

    @staticmethod
    def backward(ctx, g_colors, g_opacity, g_depth, g_weights, g_trans, g_alphas):
        field, rgbs, t_starts, t_ends, trans, packed_info, ray_indices = ctx.saved_tensors[:7]
        opacity, depth = ctx.saved_tensors[7:] if ctx.norm_depth else (None, None)
        dev = field.device
        n = field.numel()
        need_field, need_rgb = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        g_field = torch.empty_like(field)
        g_rgbs = torch.empty_like(rgbs) if need_rgb else None
        if n > 0 and ctx.norm_depth and g_depth is not None:
            # gradient through the quotient: replaces (g_opacity, g_depth) by the gradients of the sums
            g_o0 = torch.empty_like(opacity)
            g_d0 = torch.empty_like(depth)
            with torch.cuda.device(dev):
                rc = _C.lib().nfb_depth_normalize_bwd(_p(_cg(g_opacity)), _p(_cg(g_depth)), _p(opacity),
                                                      _p(depth), ctx.n_rays, g_o0.data_ptr(), g_d0.data_ptr(),
                                                      _C.stream_ptr(dev))
            _C.check(rc, "depth_normalize_bwd")
            g_opacity, g_depth = g_o0, g_d0
        if n > 0:
            fd = ctx.from_density
            with torch.cuda.device(dev):
                if ctx.mode == "bits":
                    bits = ctx.bits
                    ws, ws_bytes, epoch = _C.workspace(dev, n)
                    rc = _C.lib().nfb_render_bits_bwd(
                        int(fd), _p(_cg(g_colors)), _p(_cg(g_opacity)), _p(_cg(g_depth)), _p(_cg(g_weights)),
                        _p(_cg(g_trans)), _p(_cg(g_alphas)) if fd else None,
                        None if fd else _p(field), _p(t_starts), _p(t_ends), _p(field) if fd else None,
                        _p(rgbs), _p(trans), _p(ray_indices), _p(bits), n,
                        g_field.data_ptr(), _p(g_rgbs), ws.data_ptr(), ws_bytes, epoch, _C.stream_ptr(dev))
                elif ctx.per_ray:
                    rc = _C.lib().nfb_render_rays_bwd(
                        int(fd), _p(_cg(g_colors)), _p(_cg(g_opacity)), _p(_cg(g_depth)), _p(_cg(g_weights)),
                        _p(_cg(g_trans)), _p(_cg(g_alphas)) if fd else None,
                        None if fd else _p(field), _p(t_starts), _p(t_ends), _p(field) if fd else None,
                        _p(rgbs), _p(trans), _p(packed_info), ctx.n_rays, n,
                        g_field.data_ptr(), _p(g_rgbs), _RAYS_L, _C.stream_ptr(dev))
                else:
                    ws, ws_bytes, epoch = _C.workspace(dev, n)
                    rc = _C.lib().nfb_composite_bwd(
                        int(fd), _p(_cg(g_colors)), _p(_cg(g_opacity)), _p(_cg(g_depth)), _p(_cg(g_weights)),
                        _p(_cg(g_trans)), _p(_cg(g_alphas)) if fd else None,
                        None if fd else _p(field), _p(t_starts), _p(t_ends), _p(field) if fd else None,
                        _p(rgbs), _p(trans), _p(ray_indices), _p(packed_info), ctx.n_rays, n,
                        g_field.data_ptr(), _p(g_rgbs), ws.data_ptr(), ws_bytes, epoch, _C.stream_ptr(dev))
            _C.check(rc, "composite_bwd")
        return (g_field if need_field else None), g_rgbs, None, None, None, None, None, None, None
