"""Volume rendering — drop-in for nerfacc/volrend.py (same names, signatures, asserts, return
conventions and autograd behaviour; reference nerfacc/volrend.py:14-573).

Where the reference composes each function from 5-7 elementwise torch kernels around one
packed scan, CUDA float32 inputs here take ONE fused kernel per direction
(nfb_render_weight_fwd / _bwd, nfb_accumulate_fwd / _bwd in include/nerfacc_b200.h).
Everything the fused path does not cover (CPU tensors, other dtypes, a hand-made
non-canonical packed_info, gradients wrt t_starts / t_ends / prefix_trans) goes through the
reference's own composite formulation on top of this package's scans, so results and
gradients are defined everywhere the reference defines them.
"""
from typing import Callable, Dict, Optional, Tuple

import torch
from torch import Tensor

from . import _render_ops as _ops
from ._packing import as_packed_info, dense_total, recall_packed_info
from .pack import pack_info
from .scan import exclusive_prod, exclusive_sum


# ----------------------------------------------------------------------------- helpers
def _resolve_segments(like: Tensor, packed_info: Optional[Tensor], ray_indices: Optional[Tensor],
                      n_rays: Optional[int]):
    """Returns (packed_info or None, canonical, uniform_len, fused_ok)."""
    if ray_indices is not None and packed_info is None:
        packed_info = pack_info(ray_indices, n_rays)
    fused = like.is_cuda and like.dtype == torch.float32
    if packed_info is not None:
        packed_info, canon = as_packed_info(packed_info)
        return packed_info, canon, 0, fused and canon and like.dim() == 1
    # batched [..., n] (or a single flat row): every row is one segment
    uniform_len = like.shape[-1] if like.dim() >= 1 else 1
    return None, False, uniform_len, fused and like.dim() >= 1 and uniform_len > 0


def _no_grad_needed(*ts: Optional[Tensor]) -> bool:
    return not (torch.is_grad_enabled() and any(t is not None and t.requires_grad for t in ts))


def _prep_prefix(prefix_trans: Optional[Tensor], like: Tensor) -> Optional[Tensor]:
    if prefix_trans is None:
        return None
    return prefix_trans.detach().to(like.dtype).expand_as(like).contiguous()


def rendering(
    # ray marching results
    t_starts: Tensor,
    t_ends: Tensor,
    ray_indices: Optional[Tensor] = None,
    n_rays: Optional[int] = None,
    # radiance field
    rgb_sigma_fn: Optional[Callable] = None,
    rgb_alpha_fn: Optional[Callable] = None,
    # rendering options
    render_bkgd: Optional[Tensor] = None,
) -> Tuple[Tensor, Tensor, Tensor, Dict]:
    """Render the rays through the radience field defined by `rgb_sigma_fn` / `rgb_alpha_fn`.

    Differentiable wrt the outputs of the callback; supports batched and flattened inputs (for
    flattened inputs pass `ray_indices` and `n_rays`).  Returns colors (n_rays, 3), opacities
    (n_rays, 1), depths (n_rays, 1) and extras {"weights", "trans", "alphas", "rgbs"[, "sigmas"]}.
    Reference: nerfacc/volrend.py:14-158.
    """
    flat = ray_indices is not None
    if flat:
        assert (
            t_starts.shape == t_ends.shape == ray_indices.shape
        ), "Since nerfacc 0.5.0, t_starts, t_ends and ray_indices must have the same shape (N,). "
    if rgb_sigma_fn is None and rgb_alpha_fn is None:
        raise ValueError(
            "At least one of `rgb_sigma_fn` and `rgb_alpha_fn` should be specified."
        )
    use_density = rgb_sigma_fn is not None
    field_fn = rgb_sigma_fn if use_density else rgb_alpha_fn
    what = "sigmas" if use_density else "alphas"

    # Query the radiance field (gradients start here); an empty batch skips the callback.
    if t_starts.shape[0] != 0:
        rgbs, field = field_fn(t_starts, t_ends, ray_indices)
    else:
        rgbs = torch.empty((0, 3), device=t_starts.device)
        field = torch.empty((0,), device=t_starts.device)
    assert rgbs.shape[-1] == 3, "rgbs must have 3 channels, got {}".format(rgbs.shape)
    assert field.shape == t_starts.shape, "{} must have shape of (N,)! Got {}".format(
        what, field.shape
    )

    if flat and _composite_fused_ok(t_starts, t_ends, ray_indices, n_rays, rgbs, field):
        # one weights kernel + one composite kernel forward, one kernel backward
        pi = pack_info(ray_indices, n_rays)
        # depth / max(opacity, eps) and its gradient are one launch each inside the Function
        # (no [n_rays] torch ops and autograd nodes around the two big kernels)
        fused_norm = True
        colors, opacities, depths, weights, trans, alphas = _ops.RenderComposite.apply(
            field, rgbs, t_starts.detach(), t_ends.detach(), pi, ray_indices.contiguous(), int(n_rays),
            use_density, fused_norm)
        if use_density:
            extras = {"weights": weights, "alphas": alphas, "trans": trans, "sigmas": field, "rgbs": rgbs}
        else:
            extras = {"weights": weights, "trans": trans, "rgbs": rgbs, "alphas": field}
        if not fused_norm:
            depths = depths / opacities.clamp_min(torch.finfo(rgbs.dtype).eps)
        if render_bkgd is not None:
            colors = colors + render_bkgd * (1.0 - opacities)
        return colors, opacities, depths, extras

    seg = dict(ray_indices=ray_indices, n_rays=n_rays)
    if use_density:
        weights, trans, alphas = render_weight_from_density(t_starts, t_ends, field, **seg)
        extras = {"weights": weights, "alphas": alphas, "trans": trans, "sigmas": field, "rgbs": rgbs}
    else:
        weights, trans = render_weight_from_alpha(field, **seg)
        extras = {"weights": weights, "trans": trans, "rgbs": rgbs, "alphas": field}

    # Composite colour, opacity and (opacity-normalised) depth per ray.
    t_mids = (t_starts + t_ends)[..., None] / 2.0
    colors = accumulate_along_rays(weights, values=rgbs, **seg)
    opacities = accumulate_along_rays(weights, values=None, **seg)
    depths = accumulate_along_rays(weights, values=t_mids, **seg)
    depths = depths / opacities.clamp_min(torch.finfo(rgbs.dtype).eps)
    if render_bkgd is not None:
        colors = colors + render_bkgd * (1.0 - opacities)
    return colors, opacities, depths, extras


def render_transmittance_from_alpha(
    alphas: Tensor,
    packed_info: Optional[Tensor] = None,
    ray_indices: Optional[Tensor] = None,
    n_rays: Optional[int] = None,
    prefix_trans: Optional[Tensor] = None,
) -> Tensor:
    """T_i = prod_{j<i} (1 - alpha_j) [* prefix_trans], per ray (batched rows or packed chunks).

    >>> alphas = torch.tensor([0.4, 0.8, 0.1, 0.8, 0.1, 0.0, 0.9], device="cuda")
    >>> ray_indices = torch.tensor([0, 0, 0, 1, 1, 2, 2], device="cuda")
    >>> render_transmittance_from_alpha(alphas, ray_indices=ray_indices)
    tensor([1.0, 0.6, 0.12, 1.0, 0.2, 1.0, 1.0])

    Reference: nerfacc/volrend.py:161-206.
    """
    _, trans = _weights_from_alpha(alphas, packed_info, ray_indices, n_rays, prefix_trans, False)
    return trans


def render_transmittance_from_density(
    t_starts: Tensor,
    t_ends: Tensor,
    sigmas: Tensor,
    packed_info: Optional[Tensor] = None,
    ray_indices: Optional[Tensor] = None,
    n_rays: Optional[int] = None,
    prefix_trans: Optional[Tensor] = None,
) -> Tuple[Tensor, Tensor]:
    """T_i = exp(-sum_{j<i} sigma_j * delta_j) [* prefix_trans] and alpha_i = 1 - exp(-sigma_i delta_i).

    Returns (transmittance, alphas), both shaped like `sigmas`.
    Reference: nerfacc/volrend.py:209-264.
    """
    _, trans, alphas = _weights_from_density(
        t_starts, t_ends, sigmas, packed_info, ray_indices, n_rays, prefix_trans, False
    )
    return trans, alphas


def render_weight_from_alpha(
    alphas: Tensor,
    packed_info: Optional[Tensor] = None,
    ray_indices: Optional[Tensor] = None,
    n_rays: Optional[int] = None,
    prefix_trans: Optional[Tensor] = None,
) -> Tuple[Tensor, Tensor]:
    """w_i = T_i * alpha_i with T_i = prod_{j<i} (1 - alpha_j).  Returns (weights, transmittance).

    >>> alphas = torch.tensor([0.4, 0.8, 0.1, 0.8, 0.1, 0.0, 0.9], device="cuda")
    >>> ray_indices = torch.tensor([0, 0, 0, 1, 1, 2, 2], device="cuda")
    >>> render_weight_from_alpha(alphas, ray_indices=ray_indices)
    weights: [0.4, 0.48, 0.012, 0.8, 0.02, 0.0, 0.9], transmittance: [1.0, 0.6, 0.12, 1.0, 0.2, 1.0, 1.0]

    Reference: nerfacc/volrend.py:267-309.
    """
    return _weights_from_alpha(alphas, packed_info, ray_indices, n_rays, prefix_trans, True)


def render_weight_from_density(
    t_starts: Tensor,
    t_ends: Tensor,
    sigmas: Tensor,
    packed_info: Optional[Tensor] = None,
    ray_indices: Optional[Tensor] = None,
    n_rays: Optional[int] = None,
    prefix_trans: Optional[Tensor] = None,
) -> Tuple[Tensor, Tensor, Tensor]:
    """w_i = T_i (1 - exp(-sigma_i delta_i)).  Returns (weights, transmittance, alphas).

    Reference: nerfacc/volrend.py:312-362.
    """
    return _weights_from_density(
        t_starts, t_ends, sigmas, packed_info, ray_indices, n_rays, prefix_trans, True
    )


def _weights_from_alpha(alphas, packed_info, ray_indices, n_rays, prefix_trans, want_weights):
    pi, canon, uniform_len, fused = _resolve_segments(alphas, packed_info, ray_indices, n_rays)
    if fused and _no_grad_needed(prefix_trans):
        w, trans = _ops.WeightFromAlpha.apply(
            alphas, _prep_prefix(prefix_trans, alphas), pi, uniform_len, want_weights
        )
        return (w if want_weights else None), trans
    # composite formulation of the reference (volrend.py:203-205, 308)
    trans = exclusive_prod(1 - alphas, pi if packed_info is None else packed_info)
    if prefix_trans is not None:
        trans = trans * prefix_trans
    return (trans * alphas if want_weights else None), trans


def _weights_from_density(t_starts, t_ends, sigmas, packed_info, ray_indices, n_rays,
                          prefix_trans, want_weights):
    pi, canon, uniform_len, fused = _resolve_segments(sigmas, packed_info, ray_indices, n_rays)
    same = t_starts.shape == sigmas.shape and t_ends.shape == sigmas.shape
    if (fused and same and t_starts.dtype == torch.float32 and t_ends.dtype == torch.float32
            and _no_grad_needed(t_starts, t_ends, prefix_trans)):
        w, trans, alphas = _ops.WeightFromDensity.apply(
            sigmas, t_starts.detach(), t_ends.detach(), _prep_prefix(prefix_trans, sigmas), pi,
            uniform_len, want_weights
        )
        return (w if want_weights else None), trans, alphas
    # composite formulation of the reference (volrend.py:259-263, 361)
    sigmas_dt = sigmas * (t_ends - t_starts)
    alphas = 1.0 - torch.exp(-sigmas_dt)
    trans = torch.exp(-exclusive_sum(sigmas_dt, pi if packed_info is None else packed_info))
    if prefix_trans is not None:
        trans = trans * prefix_trans
    return (trans * alphas if want_weights else None), trans, alphas


@torch.no_grad()
def render_visibility_from_alpha(
    alphas: Tensor,
    packed_info: Optional[Tensor] = None,
    ray_indices: Optional[Tensor] = None,
    n_rays: Optional[int] = None,
    early_stop_eps: float = 1e-4,
    alpha_thre: float = 0.0,
    prefix_trans: Optional[Tensor] = None,
) -> Tensor:
    """Boolean mask of samples with transmittance >= early_stop_eps (and alpha >= alpha_thre
    when alpha_thre > 0).

    >>> alphas = torch.tensor([0.4, 0.8, 0.1, 0.8, 0.1, 0.0, 0.9], device="cuda")
    >>> ray_indices = torch.tensor([0, 0, 0, 1, 1, 2, 2], device="cuda")
    >>> render_visibility_from_alpha(alphas, ray_indices=ray_indices, early_stop_eps=0.3, alpha_thre=0.2)
    tensor([True,  True, False,  True, False, False,  True])

    Reference: nerfacc/volrend.py:365-418.
    """
    pi, canon, uniform_len, fused = _resolve_segments(alphas, packed_info, ray_indices, n_rays)
    if fused:
        _, _, _, vis = _ops.weight_fwd(
            False, alphas.contiguous(), None, None, None, _prep_prefix(prefix_trans, alphas), pi,
            uniform_len, False, False, False, vis=(early_stop_eps, alpha_thre))
        return vis
    _, trans = _weights_from_alpha(alphas, packed_info, ray_indices, n_rays, prefix_trans, False)
    vis = trans >= early_stop_eps
    if alpha_thre > 0:
        vis = vis & (alphas >= alpha_thre)
    return vis


@torch.no_grad()
def render_visibility_from_density(
    t_starts: Tensor,
    t_ends: Tensor,
    sigmas: Tensor,
    packed_info: Optional[Tensor] = None,
    ray_indices: Optional[Tensor] = None,
    n_rays: Optional[int] = None,
    early_stop_eps: float = 1e-4,
    alpha_thre: float = 0.0,
    prefix_trans: Optional[Tensor] = None,
) -> Tensor:
    """Same as :func:`render_visibility_from_alpha` with alpha = 1 - exp(-sigma * (t_end - t_start)).

    Reference: nerfacc/volrend.py:421-480.
    """
    pi, canon, uniform_len, fused = _resolve_segments(sigmas, packed_info, ray_indices, n_rays)
    if (fused and t_starts.shape == sigmas.shape == t_ends.shape
            and t_starts.dtype == torch.float32 and t_ends.dtype == torch.float32):
        _, _, _, vis = _ops.weight_fwd(
            True, None, t_starts.contiguous(), t_ends.contiguous(), sigmas.contiguous(),
            _prep_prefix(prefix_trans, sigmas), pi, uniform_len, False, False, False,
            vis=(early_stop_eps, alpha_thre))
        return vis
    _, trans, alphas = _weights_from_density(
        t_starts, t_ends, sigmas, packed_info, ray_indices, n_rays, prefix_trans, False
    )
    vis = trans >= early_stop_eps
    if alpha_thre > 0:
        vis = vis & (alphas >= alpha_thre)
    return vis


def accumulate_along_rays(
    weights: Tensor,
    values: Optional[Tensor] = None,
    ray_indices: Optional[Tensor] = None,
    n_rays: Optional[int] = None,
) -> Tensor:
    """Accumulate volumetric values along the ray: out[r] = sum_i weights[i] * values[i].

    Batched inputs (n_rays, n_samples[, D]) reduce over the sample dim; flattened inputs
    (all_samples[, D]) need `ray_indices` and `n_rays`.  `values=None` accumulates the weights
    themselves (D == 1).  Differentiable wrt `weights` and `values`.  Returns (n_rays, D).

    Reference: nerfacc/volrend.py:483-547 (temporary product + index_add_ atomics there; a
    segmented reduction here).
    """
    if values is not None:
        assert values.dim() == weights.dim() + 1
        assert weights.shape == values.shape[:-1]
    if ray_indices is None:
        src = weights[..., None] if values is None else weights[..., None] * values
        return torch.sum(src, dim=-2)
    assert n_rays is not None, "n_rays must be provided"
    assert weights.dim() == 1, "weights must be flattened"
    assert ray_indices.shape == weights.shape, (
        f"ray_indices {tuple(ray_indices.shape)} and weights {tuple(weights.shape)} must have the same shape")
    if _accumulate_fused_ok(weights, values, ray_indices):
        ri = ray_indices.contiguous()
        pi = _recall_for(ray_indices, n_rays, weights.numel())
        return _ops.Accumulate.apply(weights, values, pi, ri, int(n_rays))
    src = weights[..., None] if values is None else weights[..., None] * values
    outputs = torch.zeros((n_rays, src.shape[-1]), device=src.device, dtype=src.dtype)
    outputs.index_add_(0, ray_indices, src)
    return outputs


def accumulate_along_rays_(
    weights: Tensor,
    values: Optional[Tensor] = None,
    ray_indices: Optional[Tensor] = None,
    outputs: Optional[Tensor] = None,
) -> None:
    """In-place version of :func:`accumulate_along_rays`: `outputs` (n_rays, D) += the ray sums.

    Reference: nerfacc/volrend.py:550-573.
    """
    if values is not None:
        assert values.dim() == weights.dim() + 1
        assert weights.shape == values.shape[:-1]
    dim = 1 if values is None else values.shape[-1]
    if ray_indices is None:
        src = weights[..., None] if values is None else weights[..., None] * values
        outputs.add_(src.sum(dim=-2))
        return
    assert weights.dim() == 1, "weights must be flattened"
    assert (
        outputs.dim() == 2 and outputs.shape[-1] == dim
    ), "outputs must be of shape (n_rays, D)"
    assert ray_indices.shape == weights.shape, (
        f"ray_indices {tuple(ray_indices.shape)} and weights {tuple(weights.shape)} must have the same shape")
    if (_accumulate_fused_ok(weights, values, ray_indices) and outputs.is_contiguous()
            and outputs.dtype == torch.float32 and _no_grad_needed(weights, values, outputs)):
        pi = _recall_for(ray_indices, outputs.shape[0], weights.numel())
        _ops.accumulate_fwd(weights.detach().contiguous(),
                            None if values is None else values.detach().contiguous(), pi,
                            ray_indices.contiguous(), outputs.shape[0], out=outputs)
        return
    src = weights[..., None] if values is None else weights[..., None] * values
    outputs.index_add_(0, ray_indices, src)


def _recall_for(ray_indices: Tensor, n_rays: int, n: int) -> Optional[Tensor]:
    """The packed_info remembered for `ray_indices`, only if it describes exactly n samples."""
    pi = recall_packed_info(ray_indices, n_rays)
    if pi is not None and dense_total(pi) != int(n):
        return None
    return pi


def _composite_fused_ok(t_starts, t_ends, ray_indices, n_rays, rgbs, field) -> bool:
    f32 = torch.float32
    return (
        n_rays is not None
        and field.is_cuda and field.dim() == 1 and field.numel() > 0
        and field.dtype == f32 and rgbs.dtype == f32 and t_starts.dtype == f32 and t_ends.dtype == f32
        and rgbs.dim() == 2 and rgbs.shape[0] == field.shape[0]
        and ray_indices.dtype == torch.long and ray_indices.is_cuda
        and _no_grad_needed(t_starts, t_ends)
    )


def _accumulate_fused_ok(weights, values, ray_indices) -> bool:
    return (
        weights.is_cuda
        and weights.dtype == torch.float32
        and ray_indices.dtype == torch.long
        and (values is None or (values.dtype == torch.float32 and values.shape[-1] > 0))
    )
