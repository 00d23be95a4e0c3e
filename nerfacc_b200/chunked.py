"""The test-time rendering loop of nerfacc's callers as a library call (additive API).

The reference ships this loop in its examples (examples/utils.py:252-425,
`render_image_with_occgrid_test`): march a few samples per ray, query the radiance field,
composite behind what the rays already hold, retire rays that are opaque or have left the grid,
repeat.  Spelled with the public API it costs, per iteration, `ray_mask.sum().item()`,
`traverse_grids(over_allocate=True, ...)` (4 kernels + two cumsums), three boolean-mask selects
(each a sync), `render_weight_from_density`, optional mask filtering, three
`accumulate_along_rays_` and the mask update — a dozen launches and 4-5 host round trips around a
few thousand samples.  Here the loop state lives on the device (csrc/grid.cu, "chunked loop"): an
iteration is ONE march launch group, the user's field, ONE composite launch group, and one host
read (the chunk's sample total, which sizes the tensors the field sees).
"""
from __future__ import annotations

import ctypes as C
from typing import Callable, Optional, Tuple

import torch
from torch import Tensor

from . import _C
from ._grid_ops import Marcher
from .grid import ray_aabb_intersect


@torch.no_grad()
def render_image_with_occgrid_test(
    max_samples: int,
    rgb_sigma_fn: Callable,
    estimator,
    rays_o: Tensor,
    rays_d: Tensor,
    near_plane: float = 0.0,
    far_plane: float = 1e10,
    render_step_size: float = 1e-3,
    render_bkgd: Optional[Tensor] = None,
    cone_angle: float = 0.0,
    alpha_thre: float = 0.0,
    early_stop_eps: float = 1e-4,
    _on_chunk: Optional[Callable] = None,
) -> Tuple[Tensor, Tensor, Tensor, int]:
    """Render rays by chunked marching with early termination — the semantics, argument names and
    return values of the reference's `render_image_with_occgrid_test` (examples/utils.py:252-425),
    with the radiance field given as `rgb_sigma_fn(t_starts, t_ends, ray_indices) -> (rgbs, sigmas)`.

    Args:
        max_samples: bound on the samples marched per ray (sum of the per-iteration chunk sizes).
        rgb_sigma_fn: the field; called once per iteration on the chunk's samples (flattened).
        estimator: an `OccGridEstimator` (its `binaries` and `aabbs` are used).
        rays_o, rays_d: (n_rays, 3) or (H, W, 3).
        near_plane, far_plane, render_step_size, cone_angle: as in `OccGridEstimator.sampling`.
        render_bkgd: optional background colour blended behind the rays.
        alpha_thre: samples with alpha below it do not contribute.
        early_stop_eps: a ray retires once its opacity exceeds 1 - early_stop_eps.

    Returns:
        rgb (..., 3), opacity (..., 1), depth (..., 1) and the total number of samples marched.
    """
    shape = rays_o.shape
    o = rays_o.reshape(-1, 3).float().contiguous()
    d = rays_d.reshape(-1, 3).float().contiguous()
    dev = _C.require_cuda(o, d, estimator.binaries, estimator.aabbs)
    n = o.shape[0]
    rgb = torch.zeros((n, 3), device=dev)
    opacity = torch.zeros((n, 1), device=dev)
    depth = torch.zeros((n, 1), device=dev)
    total_samples = 0
    if n > 0:
        min_samples = 1 if cone_angle == 0 else 4  # 1 for synthetic scenes, 4 for real scenes (utils.py:312)
        near = torch.full((n,), float(near_plane), device=dev)
        far = torch.full((n,), float(far_plane), device=dev)
        m = estimator.binaries.shape[0]
        pre = {}
        if m > 1:  # the same events for every chunk: intersect and sort once (utils.py:322-331)
            t_mins, t_maxs, hits = ray_aabb_intersect(o, d, estimator.aabbs)
            t_sorted, t_indices = torch.sort(torch.cat([t_mins, t_maxs], -1), -1)
            pre = dict(t_sorted=t_sorted, t_indices=t_indices, hits=hits)
        marcher = Marcher(o, d, estimator.binaries, estimator.aabbs, near, far, render_step_size, cone_angle, 64, **pre)
        lib = _C.lib()
        cap = n * max(min_samples, 1)
        state = torch.zeros(int(lib.nfb_chunk_state_bytes()), dtype=torch.uint8, device=dev)
        alive = [torch.arange(n, dtype=torch.int32, device=dev), torch.empty(n, dtype=torch.int32, device=dev)]
        slot_of = torch.arange(n, dtype=torch.long, device=dev)
        sample_info = torch.zeros((n, 2), dtype=torch.long, device=dev)
        slot_ts, slot_te = torch.empty(cap, device=dev), torch.empty(cap, device=dev)
        ri = torch.empty(cap, dtype=torch.long, device=dev)
        ts, te = torch.empty(cap, device=dev), torch.empty(cap, device=dev)
        host = torch.zeros(8, dtype=torch.long).pin_memory()  # [0:3] totals of the march, [4:6] the plan
        totals, plan = host[0:3], host[4:6]
        opc_thre = 1 - early_stop_eps
        with torch.cuda.device(dev):
            stream = _C.stream_ptr(dev)
            sc = _C.scratch(dev, int(lib.nfb_traverse_scratch_bytes(n)), "ray")
            _C.check(lib.nfb_chunk_init(state.data_ptr(), n, min_samples, plan.data_ptr(), stream), "chunk_init")
            n_alive_max, iter_samples, cur = n, 0, 0
            while iter_samples < max_samples:
                near_before = near.clone() if _on_chunk is not None else None
                _C.check(lib.nfb_chunk_march(
                    C.byref(marcher.rays), C.byref(marcher.grid), C.byref(marcher.march), alive[cur].data_ptr(),
                    n_alive_max, state.data_ptr(), slot_of.data_ptr(), sample_info.data_ptr(), slot_ts.data_ptr(),
                    slot_te.data_ptr(), ri.data_ptr(), ts.data_ptr(), te.data_ptr(), totals.data_ptr(),
                    sc.data_ptr(), sc.numel(), stream), "chunk_march")
                torch.cuda.current_stream(dev).synchronize()  # the iteration's one host round trip
                n_alive, k, S = int(plan[0]), int(plan[1]), int(totals[0])
                if n_alive == 0:
                    break  # (the march above found an empty list and did nothing)
                iter_samples += k
                total_samples += S
                if _on_chunk is not None:
                    _on_chunk(dict(k=k, alive=alive[cur][:n_alive], near_before=near_before, near_after=near,
                                   packed_info=sample_info, ray_indices=ri[:S], t_starts=ts[:S], t_ends=te[:S]))
                if S == 0:
                    break  # rays were alive but none found another sample: all of them have left the grid
                rgbs, sigmas = rgb_sigma_fn(ts[:S], te[:S], ri[:S])
                rgbs, sigmas = rgbs.float().contiguous(), sigmas.float().contiguous().reshape(-1)
                assert rgbs.shape == (S, 3) and sigmas.shape == (S,), (rgbs.shape, sigmas.shape)
                _C.check(lib.nfb_chunk_composite(
                    alive[cur].data_ptr(), alive[1 - cur].data_ptr(), n_alive, state.data_ptr(),
                    sample_info.data_ptr(), slot_of.data_ptr(), ts.data_ptr(), te.data_ptr(), sigmas.data_ptr(),
                    rgbs.data_ptr(), n, rgb.data_ptr(), opacity.data_ptr(), depth.data_ptr(), float(alpha_thre),
                    float(opc_thre), min_samples, plan.data_ptr(), stream), "chunk_composite")
                cur = 1 - cur
                n_alive_max = n_alive  # the next list is no longer than this one
    if render_bkgd is not None:
        rgb = rgb + render_bkgd * (1.0 - opacity)
    depth = depth / opacity.clamp_min(torch.finfo(torch.float32).eps)
    lead = tuple(shape[:-1])
    return rgb.view(lead + (3,)), opacity.view(lead + (1,)), depth.view(lead + (1,)), total_samples
