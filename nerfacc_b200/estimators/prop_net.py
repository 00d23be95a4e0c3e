"""PropNetEstimator — proposal-network transmittance estimator (Mip-NeRF 360), drop-in for
nerfacc/estimators/prop_net.py:22-193 (same constructor, `sampling`, `compute_loss`,
`update_every_n_steps`, `prop_cache` protocol) plus its helpers `_transform_stot`, `_pdf_loss`,
`get_proposal_requires_grad_fn` (prop_net.py:195-259).

Device work per level: one `nfb_importance_sampling_stot` launch (the edges in s and, mapped
through `_transform_stot`, in t; the centres are not needed and not written), the user's proposal
network, and one `nfb_cdf_from_density` launch (interval lengths, exclusive sum, exp, `1 - cat([T, 0])`
in one pass; its backward is one launch too).  The loss is one `nfb_pdf_loss` launch per cached
level each way (searchsorted, envelope mass, penalty; the gradient row of `cdfs_key` without atomics).
Rows that the fused kernels do not take (packed intervals, tensor-valued near/far, rows of more
than 512 samples) run the reference's sequence of ops on the general kernels.
"""
from typing import Callable, List, Optional, Tuple

try:
    from typing import Literal
except ImportError:  # pragma: no cover
    from typing_extensions import Literal

import torch
from torch import Tensor

from .. import _prop_ops
from ..data_specs import RayIntervals
from ..pdf import importance_sampling, searchsorted
from ..volrend import render_transmittance_from_density
from .base import AbstractEstimator


def _cdf_from_trans(trans: Tensor) -> Tensor:
    """CDF at the n+1 edges of n samples: 1 - T, closed with a final 1 (prop_net.py:117-119)."""
    return 1.0 - torch.cat([trans, torch.zeros_like(trans[:, :1])], dim=-1)


class PropNetEstimator(AbstractEstimator):
    """Proposal network transmittance estimator.

    Args:
        optimizer: optimizer of the proposal networks (needed only for `update_every_n_steps`
            with `requires_grad=True`).
        scheduler: optional learning-rate scheduler, stepped once per update call.
    """

    def __init__(
        self,
        optimizer: Optional[torch.optim.Optimizer] = None,
        scheduler: Optional[torch.optim.lr_scheduler._LRScheduler] = None,
    ) -> None:
        super().__init__()
        self.optimizer = optimizer
        self.scheduler = scheduler
        self.prop_cache: List = []

    @torch.no_grad()
    def sampling(
        self,
        prop_sigma_fns: List[Callable],
        prop_samples: List[int],
        num_samples: int,
        n_rays: int,
        near_plane: float,
        far_plane: float,
        sampling_type: Literal["uniform", "lindisp"] = "lindisp",
        stratified: bool = False,
        requires_grad: bool = False,
    ) -> Tuple[Tensor, Tensor]:
        """Hierarchical sampling through the proposal networks (prop_net.py:45-136).

        Starts from the uniform CDF on s in [0,1]; every level resamples `prop_samples[i]`
        intervals from the current CDF, maps them to t with `sampling_type`, evaluates
        `prop_sigma_fns[i](t_starts, t_ends) -> sigmas` (all (n_rays, n)) and turns the densities
        into the next CDF.  A last resampling gives `num_samples` intervals.  With
        `requires_grad` the level outputs are kept in `prop_cache` (and the networks run with
        autograd on) for `compute_loss` / `update_every_n_steps`.

        Returns `t_starts`, `t_ends`, both (n_rays, num_samples).
        """
        assert len(prop_sigma_fns) == len(prop_samples), (
            "The number of proposal networks and the number of samples should be the same."
        )
        dev = self.device
        cdfs = torch.tensor([0.0, 1.0], device=dev).repeat(n_rays, 1)
        intervals = RayIntervals(vals=cdfs)

        stot = (sampling_type, near_plane, far_plane)
        if sampling_type not in ("uniform", "lindisp"):
            raise ValueError(f"Unknown transform_type: {sampling_type}")
        for level_fn, level_samples in zip(prop_sigma_fns, prop_samples):
            intervals, _, t_vals = importance_sampling(intervals, cdfs, level_samples, stratified,
                                                       _samples=False, _stot=stot)
            if t_vals is None:
                t_vals = _transform_stot(sampling_type, intervals.vals, near_plane, far_plane)
            t_starts, t_ends = t_vals[..., :-1], t_vals[..., 1:]
            with torch.set_grad_enabled(requires_grad):
                sigmas = level_fn(t_starts, t_ends)
                assert sigmas.shape == t_starts.shape
                if _prop_ops.cdf_rows_ok(t_vals, sigmas):
                    cdfs = _prop_ops.CdfFromDensity.apply(t_vals, sigmas)
                else:
                    trans, _ = render_transmittance_from_density(t_starts, t_ends, sigmas)
                    cdfs = _cdf_from_trans(trans)
                if requires_grad:
                    self.prop_cache.append((intervals, cdfs))

        intervals, _, t_vals = importance_sampling(intervals, cdfs, num_samples, stratified,
                                                   _samples=False, _stot=stot)
        if t_vals is None:
            t_vals = _transform_stot(sampling_type, intervals.vals, near_plane, far_plane)
        if requires_grad:
            self.prop_cache.append((intervals, None))
        return t_vals[..., :-1], t_vals[..., 1:]

    @torch.enable_grad()
    def compute_loss(self, trans: Tensor, loss_scaler: float = 1.0) -> Tensor:
        """Proposal loss against the final samples' transmittance `trans` (n_rays, num_samples);
        empties `prop_cache` (prop_net.py:138-160)."""
        if len(self.prop_cache) == 0:
            return torch.zeros((), device=self.device)
        intervals, _ = self.prop_cache.pop()
        cdfs = _cdf_from_trans(trans).detach()
        loss = 0.0
        while self.prop_cache:
            prop_intervals, prop_cdfs = self.prop_cache.pop()
            loss += _pdf_loss(intervals, cdfs, prop_intervals, prop_cdfs).mean()
        return loss * loss_scaler

    @torch.enable_grad()
    def update_every_n_steps(
        self,
        trans: Tensor,
        requires_grad: bool = False,
        loss_scaler: float = 1.0,
    ) -> float:
        """One training-step update (prop_net.py:162-186): with `requires_grad` optimises the
        proposal networks on the cached levels and returns the loss value; otherwise only steps
        the scheduler and returns 0.0."""
        if requires_grad:
            return self._update(trans=trans, loss_scaler=loss_scaler)
        if self.scheduler is not None:
            self.scheduler.step()
        return 0.0

    @torch.enable_grad()
    def _update(self, trans: Tensor, loss_scaler: float = 1.0) -> float:
        assert len(self.prop_cache) > 0
        assert self.optimizer is not None, "No optimizer is provided."
        loss = self.compute_loss(trans, loss_scaler)
        self.optimizer.zero_grad()
        loss.backward()
        self.optimizer.step()
        if self.scheduler is not None:
            self.scheduler.step()
        return loss.item()


def get_proposal_requires_grad_fn(target: float = 5.0, num_steps: int = 1000) -> Callable:
    """Schedule deciding at which steps the proposal networks train: the gap between training
    steps grows linearly to `target` over `num_steps` (prop_net.py:195-211)."""
    since_last = 0

    def proposal_requires_grad_fn(step: int) -> bool:
        nonlocal since_last
        wanted_gap = min(step / num_steps, 1.0) * target
        fire = since_last > wanted_gap
        if fire:
            since_last = 0
        since_last += 1
        return fire

    return proposal_requires_grad_fn


def _transform_stot(
    transform_type: Literal["uniform", "lindisp"],
    s_vals: Tensor,
    t_min,
    t_max,
) -> Tensor:
    """Normalised distance s in [0,1] -> ray distance t, linear in t ("uniform") or in
    disparity 1/t ("lindisp") (prop_net.py:214-229)."""
    if transform_type == "uniform":
        return s_vals * t_max + (1 - s_vals) * t_min
    if transform_type == "lindisp":
        return 1 / (s_vals * (1 / t_max) + (1 - s_vals) * (1 / t_min))
    raise ValueError(f"Unknown transform_type: {transform_type}")


def _pdf_loss(
    segments_query: RayIntervals,
    cdfs_query: Tensor,
    segments_key: RayIntervals,
    cdfs_key: Tensor,
    eps: float = 1e-7,
) -> Tensor:
    """Mip-NeRF 360 proposal loss: each query interval's mass must not exceed the mass of the
    key intervals that envelope it (prop_net.py:232-257)."""
    if (segments_query.vals.dim() == 2 and segments_key.vals.dim() == 2
            and _prop_ops.pdf_loss_rows_ok(segments_query.vals, cdfs_query, segments_key.vals, cdfs_key)):
        # one launch: search, envelope mass and penalty (and one launch for the gradient)
        return _prop_ops.PdfLoss.apply(segments_query.vals, cdfs_query, segments_key.vals, cdfs_key, eps)
    ids_left, ids_right = searchsorted(segments_key, segments_query)
    if segments_query.vals.dim() > 1:
        w = cdfs_query[..., 1:] - cdfs_query[..., :-1]
        ids_left, ids_right = ids_left[..., :-1], ids_right[..., 1:]
    else:
        assert segments_query.is_left is not None and segments_query.is_right is not None
        w = cdfs_query[segments_query.is_right] - cdfs_query[segments_query.is_left]
        ids_left, ids_right = ids_left[segments_query.is_left], ids_right[segments_query.is_right]
    w_outer = cdfs_key.gather(-1, ids_right) - cdfs_key.gather(-1, ids_left)
    return torch.clip(w - w_outer, min=0) ** 2 / (w + eps)


def _outer(t0_starts: Tensor, t0_ends: Tensor, t1_starts: Tensor, t1_ends: Tensor, y1: Tensor) -> Tensor:
    """Mass of histogram (t1, y1) over the t1 bins touching each t0 interval (prop_net.py:260-293)."""
    cy1 = torch.cat([torch.zeros_like(y1[..., :1]), torch.cumsum(y1, dim=-1)], dim=-1)
    top = y1.shape[-1] - 1
    lo = torch.searchsorted(t1_starts.contiguous(), t0_starts.contiguous(), side="right") - 1
    hi = torch.searchsorted(t1_ends.contiguous(), t0_ends.contiguous(), side="right")
    lo, hi = lo.clamp(0, top), hi.clamp(0, top)
    return torch.take_along_dim(cy1[..., 1:], hi, dim=-1) - torch.take_along_dim(cy1[..., :-1], lo, dim=-1)


def _lossfun_outer(t: Tensor, w: Tensor, t_env: Tensor, w_env: Tensor) -> Tensor:
    """Penalty where histogram (t, w) exceeds its envelope (t_env, w_env) (prop_net.py:296-313)."""
    eps = torch.finfo(t.dtype).eps
    w_outer = _outer(t[..., :-1], t[..., 1:], t_env[..., :-1], t_env[..., 1:], w_env)
    return torch.clip(w - w_outer, min=0) ** 2 / (w + eps)
