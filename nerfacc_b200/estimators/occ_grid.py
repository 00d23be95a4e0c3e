"""OccGridEstimator — drop-in for nerfacc/estimators/occ_grid.py (reference :14-423).

Same constructor, buffers (`resolution`, `aabbs`, `occs`, `binaries` persistent, so reference
checkpoints load), `sampling`, `update_every_n_steps`, `mark_invisible_cells` and private
helpers.  What differs is underneath:
  * `sampling` marches with the two-pass traversal of libnerfacc_b200 (one host sync for the
    sample total; the reference has eight or more), gets (ray_indices, t_starts, t_ends)
    straight from the fill pass, and evaluates visibility with the fused kernel;
  * the EMA update uses scatter-max semantics — what the reference's author intended
    (occ_grid.py:399-402) — so it is deterministic with duplicate cell ids.  By default every
    rank updates its own grid from the global RNG, as the reference does.  After
    `shard_updates(process_group)` (opt-in) the ranks of that group split the `occ_eval_fn`
    evaluations and `occs` is MAX-all-reduced (NCCL) before binarisation, giving every rank the
    single-GPU result — the path's one collective;
  * the (48 + 16) MiB int64 `grid_coords` / `grid_indices` tables of the reference are
    derived on demand instead of being resident buffers.
"""
from typing import Callable, List, Optional, Tuple, Union

import torch
from torch import Tensor

from .. import _occ_ops
from .._grid_ops import Marcher
from .._packing import mark_canonical, remember_packed_info
from ..grid import _enlarge_aabb
from .base import AbstractEstimator


class OccGridEstimator(AbstractEstimator):
    """Occupancy grid transmittance estimator for spatial skipping.

    References: "Instant Neural Graphics Primitives."

    Args:
        roi_aabb: The axis-aligned bounding box of the region of interest. Useful for mapping
            the 3D space to the grid.
        resolution: The resolution of the grid. If an integer is given, the grid is assumed to
            be a cube. Otherwise, a list or a tensor of shape (3,) is expected. Default: 128.
        levels: The number of levels of the grid. Default: 1.
    """

    DIM: int = 3

    def __init__(
        self,
        roi_aabb: Union[List[int], Tensor],
        resolution: Union[int, List[int], Tensor] = 128,
        levels: int = 1,
        **kwargs,
    ) -> None:
        super().__init__()

        if "contraction_type" in kwargs:
            raise ValueError(
                "`contraction_type` is not supported anymore for nerfacc >= 0.4.0."
            )

        # check the resolution is legal
        if isinstance(resolution, int):
            resolution = [resolution] * self.DIM
        if isinstance(resolution, (list, tuple)):
            resolution = torch.tensor(resolution, dtype=torch.int32)
        assert isinstance(resolution, Tensor), f"Invalid type: {resolution}!"
        assert resolution.shape[0] == self.DIM, f"Invalid shape: {resolution}!"

        # check the roi_aabb is legal
        if isinstance(roi_aabb, (list, tuple)):
            roi_aabb = torch.tensor(roi_aabb, dtype=torch.float32)
        assert isinstance(roi_aabb, Tensor), f"Invalid type: {roi_aabb}!"
        assert roi_aabb.shape[0] == self.DIM * 2, f"Invalid shape: {roi_aabb}!"

        # level i covers the roi scaled 2^i about its centre
        aabbs = torch.stack([_enlarge_aabb(roi_aabb, 2**i) for i in range(levels)], dim=0)

        self.cells_per_lvl = int(resolution.prod().item())
        self.levels = levels

        self.register_buffer("resolution", resolution)  # [3]
        self.register_buffer("aabbs", aabbs)  # [n_aabbs, 6]
        self.register_buffer("occs", torch.zeros(self.levels * self.cells_per_lvl))
        self.register_buffer(
            "binaries", torch.zeros([levels] + resolution.tolist(), dtype=torch.bool)
        )
        self._occ_mean_cache: Optional[Tuple[int, int, float]] = None
        self._count_hint: dict = {}  # n_rays -> samples marched by the last sampling() call of that size
        # sharded updates (opt-in, see shard_updates): (process group | None, world, rank, seed)
        self._shard: Optional[Tuple[object, int, int, int]] = None
        self._update_count = 0

    def shard_updates(self, process_group=None, seed: int = 0, enabled: bool = True,
                      _world_rank: Optional[Tuple[int, int]] = None) -> "OccGridEstimator":
        """Opt in to (or out of) splitting `_update` across the ranks of `process_group` (None =
        the default group of torch.distributed, which must be initialised).

        No counterpart in the reference, whose ranks each update their own grid.  When enabled,
        every `_update` is a COLLECTIVE: all ranks of the group must call it in lockstep with the
        same `step` and an `occ_eval_fn` of the same model, else the all-reduce deadlocks or mixes
        grids.  The cells and jitter offsets of update number k are drawn from a private generator
        seeded with (seed, k) — identical on every rank, independent of torch.manual_seed — each
        rank evaluates `occ_eval_fn` on its 1/world slice of them, applies the decay to all of
        them, and `occs` is MAX-all-reduced: the result is bit-identical to the same update done
        by one rank (SURVEY.md §8e exactness rule).  Default (never called): rank-local updates
        from the global RNG, no collective."""
        if not enabled:
            self._shard = None
            return self
        if _world_rank is not None:  # tests / parity checks: the same draws on a single rank
            world, rank = _world_rank
        else:
            world, rank = _occ_ops.dist_info(process_group)
        self._shard = (process_group, int(world), int(rank), int(seed))
        self._update_count = 0
        return self

    # -- derived tables (resident int64 buffers in the reference, occ_grid.py:76-83) ----------
    @property
    def grid_indices(self) -> Tensor:
        return torch.arange(self.cells_per_lvl, device=self.device)

    @property
    def grid_coords(self) -> Tensor:
        return _meshgrid3d(self.resolution, self.device).reshape(self.cells_per_lvl, self.DIM)

    @torch.no_grad()
    def sampling(
        self,
        # rays
        rays_o: Tensor,  # [n_rays, 3]
        rays_d: Tensor,  # [n_rays, 3]
        # sigma/alpha function for skipping invisible space
        sigma_fn: Optional[Callable] = None,
        alpha_fn: Optional[Callable] = None,
        near_plane: float = 0.0,
        far_plane: float = 1e10,
        t_min: Optional[Tensor] = None,  # [n_rays]
        t_max: Optional[Tensor] = None,  # [n_rays]
        # rendering options
        render_step_size: float = 1e-3,
        early_stop_eps: float = 1e-4,
        alpha_thre: float = 0.0,
        stratified: bool = False,
        cone_angle: float = 0.0,
    ) -> Tuple[Tensor, Tensor, Tensor]:
        """Sampling with spatial skipping.  Not differentiable.

        Args:
            rays_o, rays_d: (n_rays, 3) origins and normalized directions.
            sigma_fn / alpha_fn: optional callbacks (t_starts, t_ends, ray_indices) -> density /
                opacity (N,), used to drop invisible samples; give at most one.
            near_plane, far_plane: scalar clip range; t_min, t_max: optional per-ray (n_rays,)
                ranges, combined as max(near_plane, t_min) and min(far_plane, t_max).
            render_step_size: marching step.  early_stop_eps / alpha_thre: visibility
                thresholds on transmittance / opacity.  stratified: jitter the start by
                U[0,1) * step per ray.  cone_angle: > 0 grows the step linearly with t.

        Returns:
            ray_indices (n_samples,) long, t_starts (n_samples,), t_ends (n_samples,).

        >>> ray_indices, t_starts, t_ends = grid.sampling(rays_o, rays_d, render_step_size=1e-3)
        >>> t_mid = (t_starts + t_ends) / 2.0
        >>> sample_locs = rays_o[ray_indices] + t_mid * rays_d[ray_indices]

        Reference: nerfacc/estimators/occ_grid.py:86-221.
        """
        near, far = near_plane, far_plane
        if t_min is not None:
            near = torch.clamp(torch.full_like(rays_o[..., 0], near_plane), min=t_min)
        if t_max is not None:
            far = torch.clamp(torch.full_like(rays_o[..., 0], far_plane), max=t_max)
        if stratified:
            if not isinstance(near, Tensor):
                near = torch.full_like(rays_o[..., 0], near_plane)
            near = near + torch.rand_like(near) * render_step_size

        marcher = Marcher(rays_o, rays_d, self.binaries, self.aabbs, near, far,
                          render_step_size, cone_angle, -1)
        # A trainer's sample count drifts slowly from step to step: expand into buffers sized from the
        # previous call (+ 6 %) while the exact total is on its way to the host; a total that does not
        # fit is expanded again into exact buffers (results are the same either way).
        hint = self._count_hint.get(marcher.n_rays, 0)
        packed_info, _, _, n_samples, _ = marcher.count(
            False, False, want_runs=True, expand_capacity=(hint + hint // 16 + 4096) if hint > 0 else 0)
        self._count_hint[marcher.n_rays] = n_samples
        ray_indices, t_starts, t_ends = marcher.fill_sampling(packed_info, n_samples)
        mark_canonical(packed_info, total=n_samples)
        remember_packed_info(ray_indices, packed_info, marcher.n_rays)

        # skip invisible space
        if (alpha_thre > 0.0 or early_stop_eps > 0.0) and (
            sigma_fn is not None or alpha_fn is not None
        ):
            alpha_thre = min(alpha_thre, self._occs_mean())
            if sigma_fn is not None:
                values = (sigma_fn(t_starts, t_ends, ray_indices) if n_samples != 0
                          else torch.empty((0,), device=t_starts.device))
                assert (
                    values.shape == t_starts.shape
                ), "sigmas must have shape of (N,)! Got {}".format(values.shape)
            else:
                values = (alpha_fn(t_starts, t_ends, ray_indices) if n_samples != 0
                          else torch.empty((0,), device=t_starts.device))
                assert (
                    values.shape == t_starts.shape
                ), "alphas must have shape of (N,)! Got {}".format(values.shape)
            # visibility, the three selects and the survivors' packed_info in one kernel
            ray_indices, t_starts, t_ends, packed_info, _ = _occ_ops.visibility_compact(
                sigma_fn is not None, values, t_starts, t_ends, packed_info, ray_indices,
                early_stop_eps, alpha_thre)
            mark_canonical(packed_info, total=ray_indices.shape[0])
            remember_packed_info(ray_indices, packed_info, marcher.n_rays)
        return ray_indices, t_starts, t_ends

    def _occs_mean(self) -> float:
        """`self.occs.mean().item()` of occ_grid.py:183, cached until `occs` changes."""
        key = (id(self.occs), self.occs._version)
        if self._occ_mean_cache is None or self._occ_mean_cache[:2] != key:
            self._occ_mean_cache = (*key, float(self.occs.mean().item()))
        return self._occ_mean_cache[2]

    @torch.no_grad()
    def update_every_n_steps(
        self,
        step: int,
        occ_eval_fn: Callable,
        occ_thre: float = 1e-2,
        ema_decay: float = 0.95,
        warmup_steps: int = 256,
        n: int = 16,
    ) -> None:
        """Update the estimator every n steps during training.

        Args:
            step: Current training step.
            occ_eval_fn: A function that takes in sample locations :math:`(N, 3)` and
                returns the occupancy values :math:`(N, 1)` at those locations.
            occ_thre: Threshold used to binarize the occupancy grid. Default: 1e-2.
            ema_decay: The decay rate for EMA updates. Default: 0.95.
            warmup_steps: Sample all cells during the warmup stage. After the warmup
                stage we change the sampling strategy to 1/4 uniformly sampled cells
                together with 1/4 occupied cells. Default: 256.
            n: Update the grid every n steps. Default: 16.

        Reference: nerfacc/estimators/occ_grid.py:224-259.
        """
        if not self.training:
            raise RuntimeError(
                "You should only call this function only during training. "
                "Please call _update() directly if you want to update the "
                "field during inference."
            )
        if step % n == 0 and self.training:
            self._update(
                step=step,
                occ_eval_fn=occ_eval_fn,
                occ_thre=occ_thre,
                ema_decay=ema_decay,
                warmup_steps=warmup_steps,
            )

    @torch.no_grad()
    def mark_invisible_cells(
        self,
        K: Tensor,
        c2w: Tensor,
        width: int,
        height: int,
        near_plane: float = 0.0,
        chunk: int = 32**3,
    ) -> None:
        """Mark the cells that aren't covered by the cameras with density -1.
        Should only be executed once before training starts (one-off host-side torch code).

        Args:
            K: Camera intrinsics of shape (N, 3, 3) or (1, 3, 3).
            c2w: Camera to world poses of shape (N, 3, 4) or (N, 4, 4).
            width, height: image size in pixels.  near_plane: near plane distance.
            chunk: cells per batch (memory bound).

        Reference: nerfacc/estimators/occ_grid.py:263-332.
        """
        assert K.dim() == 3 and K.shape[1:] == (3, 3)
        assert c2w.dim() == 3 and (
            c2w.shape[1:] == (3, 4) or c2w.shape[1:] == (4, 4)
        )
        assert K.shape[0] == c2w.shape[0] or K.shape[0] == 1

        n_cams = c2w.shape[0]
        w2c_R = c2w[:, :3, :3].transpose(2, 1)  # (n_cams, 3, 3)
        w2c_T = -w2c_R @ c2w[:, :3, 3:]  # (n_cams, 3, 1)
        # one launch over cells x cameras (the reference: ~12 kernels + 2 matmuls per 32^3 chunk)
        cams = torch.cat([w2c_R.reshape(n_cams, 9), w2c_T.reshape(n_cams, 3),
                          K.expand(n_cams, 3, 3).reshape(n_cams, 9)], dim=-1).float().contiguous()
        _occ_ops.mark_invisible_cells(self.occs, cams.to(self.device), list(self.binaries.shape[1:]), self.levels,
                                      self.aabbs, width, height, near_plane)
        self._occ_mean_cache = None

    @torch.no_grad()
    def _get_all_cells(self) -> List[Tensor]:
        """Per level, the cells whose occupancy is not -1 (visible to some camera)."""
        occs = self.occs.view(self.levels, self.cells_per_lvl)
        idx = self.grid_indices
        return [idx[occs[lvl] >= 0.0] for lvl in range(self.levels)]

    @torch.no_grad()
    def _sample_uniform_and_occupied_cells(self, n: int, generator=None) -> List[Tensor]:
        """Per level: n uniformly drawn cells (with replacement, invisible ones dropped) followed
        by the occupied cells (all of them, or n drawn with replacement if there are more)."""
        out = []
        occs = self.occs.view(self.levels, self.cells_per_lvl)
        for lvl in range(self.levels):
            uniform = torch.randint(self.cells_per_lvl, (n,), device=self.device, generator=generator)
            if uniform.is_cuda:
                # both selections in two launches and one read-back, the occupied cells written
                # straight behind the uniform ones
                cells, k1, k2 = _occ_ops.select_cells(occs[lvl], uniform, self.binaries[lvl])
                if n < k2:
                    pick = torch.randint(k2, (n,), device=self.device, generator=generator)
                    out.append(torch.cat([cells[:k1], cells[k1:k1 + k2][pick]], dim=0))
                else:
                    out.append(cells[:k1 + k2])
                continue
            # (host tensors: the gloo tests of the sharded update run this method on the CPU with
            # stand-ins for the device calls)
            uniform = uniform[occs[lvl][uniform] >= 0.0]
            occupied = torch.nonzero(self.binaries[lvl].flatten())[:, 0]
            if n < len(occupied):
                pick = torch.randint(len(occupied), (n,), device=self.device, generator=generator)
                occupied = occupied[pick]
            out.append(torch.cat([uniform, occupied], dim=0))
        return out

    @torch.no_grad()
    def _update(
        self,
        step: int,
        occ_eval_fn: Callable,
        occ_thre: float = 0.01,
        ema_decay: float = 0.95,
        warmup_steps: int = 256,
    ) -> None:
        """Update the occ field in the EMA way (reference occ_grid.py:367-404)."""
        world, rank, gen, group = 1, 0, None, None
        if self._shard is not None:
            group, world, rank, seed = self._shard
            # every rank must draw the same cells and offsets: a private generator seeded alike
            gen = torch.Generator(device=self.device)
            gen.manual_seed(((0x5EED0000 + seed) << 20) + self._update_count)
            self._update_count += 1
        if step < warmup_steps:
            lvl_indices = self._get_all_cells()
        else:
            lvl_indices = self._sample_uniform_and_occupied_cells(self.cells_per_lvl // 4, gen)

        for lvl, indices in enumerate(lvl_indices):
            u = torch.rand((indices.shape[0], 3), device=self.device, generator=gen)
            cell_ids = lvl * self.cells_per_lvl + indices
            # every rank decays the full sample; only the field evaluations are sharded
            lo, hi = _occ_ops.shard_range(indices.shape[0], world, rank)
            # (the resolution as host ints: the shape of `binaries`, no device read)
            x = _occ_ops.cell_points(indices[lo:hi], u[lo:hi], list(self.binaries.shape[1:]), self.aabbs[lvl])
            occ = occ_eval_fn(x).squeeze(-1) if hi > lo else x.new_empty((0,))
            _occ_ops.ema_update(self.occs, cell_ids, occ, lo, hi, ema_decay)
        if world > 1:
            _occ_ops.all_reduce_max(self.occs, group)
        self._occ_mean_cache = None  # the kernels wrote `occs` through its data pointer
        self.binaries = _occ_ops.binarize(self.occs, occ_thre).view(self.binaries.shape)


def _meshgrid3d(res: Tensor, device: Union[torch.device, str] = "cpu") -> Tensor:
    """Integer cell coordinates [rx, ry, rz, 3] in C ("ij") order.  Reference occ_grid.py:407-423."""
    assert len(res) == 3
    rx, ry, rz = (int(v) for v in res.tolist())
    ax = [torch.arange(k, dtype=torch.long, device=device) for k in (rx, ry, rz)]
    return torch.stack(torch.meshgrid(ax, indexing="ij"), dim=-1)
