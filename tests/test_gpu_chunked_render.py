"""The test-time rendering loop of the reference's callers (examples/utils.py:252-425): march a
few samples per ray at a time with `traverse_grids(traverse_steps_limit=n, over_allocate=True,
rays_mask=alive, near_planes=termination_planes)`, composite with `prefix_trans` and
`accumulate_along_rays_`, stop rays that are opaque or have left the grid.  SURVEY.md §8(f) rank 1:
every building block of that loop is on the hot path, this drives them together.

Checks: (1) each chunk's samples are bit-identical to the reference CUDA extension's when it is
available (same call, same tensors); (2) without early stopping the chunked image matches the
one-shot `OccGridEstimator.sampling` + `rendering` image — not bit for bit: a resumed march
re-enters the grid at its termination plane, which moves a handful of samples per 10^5 in the
reference too; (3) with early stopping every ray either saturated or was marched to the end.
"""
import numpy as np
import pytest
import torch

from tests import _refcuda as R

pytestmark = pytest.mark.gpu


def _scene(device, n_rays=6000, res=64, levels=1, seed=0):
    import nerfacc_b200 as nf

    g = torch.Generator(device=device).manual_seed(seed)
    o = torch.randn((n_rays, 3), device=device, generator=g)
    o = o / o.norm(dim=-1, keepdim=True) * 3.0
    tgt = (torch.rand((n_rays, 3), device=device, generator=g) - 0.5) * 1.6
    d = tgt - o
    d = d / d.norm(dim=-1, keepdim=True)
    est = nf.OccGridEstimator([-1.0, -1.0, -1.0, 1.0, 1.0, 1.0], resolution=res, levels=levels).to(device)
    ax = (torch.arange(res, device=device) + 0.5) / res * 2 - 1
    xx, yy, zz = torch.meshgrid(ax, ax, ax, indexing="ij")
    rad = (xx * xx + yy * yy + zz * zz).sqrt()
    shell = (rad > 0.45) & (rad < 0.8) & (torch.rand(rad.shape, device=device, generator=g) < 0.7)
    bins = shell[None].repeat(levels, 1, 1, 1)
    est.binaries = bins.contiguous()
    return est, o, d


def _field(o, d):
    def rgb_sigma_fn(t_starts, t_ends, ray_indices):
        pos = o[ray_indices] + d[ray_indices] * ((t_starts + t_ends) * 0.5)[:, None]
        sigma = 6.0 * torch.exp(-((pos.norm(dim=-1) - 0.6) / 0.25) ** 2)
        return torch.sigmoid(pos * 3.0), sigma
    return rgb_sigma_fn


def _chunked(nf, est, o, d, fn, step, cone, early_stop_eps, max_samples=4096, check_ref=False):
    n = o.shape[0]
    dev = o.device
    opacity = torch.zeros((n, 1), device=dev)
    depth = torch.zeros((n, 1), device=dev)
    rgb = torch.zeros((n, 3), device=dev)
    alive = torch.ones(n, dtype=torch.bool, device=dev)
    near = torch.zeros(n, device=dev)
    far = torch.full((n,), 1e10, device=dev)
    t_mins, t_maxs, hits = nf.ray_aabb_intersect(o, d, est.aabbs)
    t_sorted, t_indices = torch.sort(torch.cat([t_mins, t_maxs], -1), -1)
    min_samples = 1 if cone == 0 else 4
    iters = total = done_iters = 0
    while iters < max_samples:
        n_alive = int(alive.sum())
        if n_alive == 0:
            break
        k = max(min(n // n_alive, 64), min_samples)
        iters += k
        done_iters += 1
        intervals, samples, term = nf.traverse_grids(
            o, d, est.binaries, est.aabbs, near, far, step, cone, k, True, alive, t_sorted, t_indices, hits)
        if check_ref and R.available():
            iv_r, sm_r, term_r = R.mod().traverse_grids(
                o, d, alive, est.binaries, est.aabbs, t_sorted, t_indices, hits, near, far,
                float(step), float(cone), True, True, True, int(k), True)
            assert torch.equal(samples.packed_info[:, 1], sm_r.chunk_cnts)
            assert torch.equal(intervals.is_left, iv_r.is_left) and torch.equal(intervals.is_right, iv_r.is_right)
            assert torch.equal(intervals.vals[intervals.is_left], iv_r.vals[iv_r.is_left])
            assert torch.equal(samples.ray_indices[samples.is_valid], sm_r.ray_indices[sm_r.is_valid])
            assert torch.equal(term[alive], term_r[alive])
        ts = intervals.vals[intervals.is_left]
        te = intervals.vals[intervals.is_right]
        ri = samples.ray_indices[samples.is_valid]
        rgbs, sigmas = fn(ts, te, ri)
        w, _, _ = nf.render_weight_from_density(ts, te, sigmas, ray_indices=ri, n_rays=n,
                                                prefix_trans=1 - opacity[ri].squeeze(-1))
        nf.accumulate_along_rays_(w, values=rgbs, ray_indices=ri, outputs=rgb)
        nf.accumulate_along_rays_(w, values=None, ray_indices=ri, outputs=opacity)
        nf.accumulate_along_rays_(w, values=(ts + te)[..., None] / 2.0, ray_indices=ri, outputs=depth)
        near = term
        alive = torch.logical_and(opacity.view(-1) <= 1 - early_stop_eps, samples.packed_info[:, 1] == k)
        total += int(ri.shape[0])
    return rgb, opacity, depth, total, done_iters


def _mostly_close(a, b, tol=2e-3, share=0.995, hard=0.08):
    """A resumed march may gain or lose a sample on a few rays (see the module docstring): all
    but a small share of the entries agree to `tol`, none is off by more than one sample's worth."""
    err = (a - b).abs()
    assert float((err <= tol + tol * b.abs()).float().mean()) >= share, float(err.max())
    assert float(err.max()) <= hard, float(err.max())


@pytest.mark.parametrize("cone,levels", [(0.0, 1), (0.004, 2)])
def test_chunked_loop_equals_one_shot(device, cone, levels):
    import nerfacc_b200 as nf

    est, o, d = _scene(device, levels=levels)
    fn = _field(o, d)
    step = 0.01
    with torch.no_grad():
        rgb, opac, depth, total, iters = _chunked(nf, est, o, d, fn, step, cone, early_stop_eps=-1.0,
                                                  check_ref=True)
        ri, ts, te = est.sampling(o, d, render_step_size=step, cone_angle=cone)
        colors1, opac1, depth1, _ = nf.rendering(ts, te, ri, o.shape[0], rgb_sigma_fn=fn)
    assert iters > 3 and ri.shape[0] > 50000 and abs(total - ri.shape[0]) <= 1e-3 * ri.shape[0]
    _mostly_close(rgb, colors1)
    _mostly_close(opac, opac1)
    _mostly_close(depth / opac.clamp_min(torch.finfo(torch.float32).eps), depth1, tol=5e-3, hard=0.3)


def test_chunked_loop_early_stop(device):
    import nerfacc_b200 as nf

    est, o, d = _scene(device, seed=3)
    fn0 = _field(o, d)
    fn = lambda a, b, r: (fn0(a, b, r)[0], fn0(a, b, r)[1] * 40.0)  # dense: rays saturate early
    with torch.no_grad():
        rgb, opac, depth, total, _ = _chunked(nf, est, o, d, fn, 0.01, 0.0, early_stop_eps=1e-3)
        ri, ts, te = est.sampling(o, d, render_step_size=0.01)
        _, opac1, _, _ = nf.rendering(ts, te, ri, o.shape[0], rgb_sigma_fn=fn)
    assert total < 0.8 * ri.shape[0]  # early stopping skipped work
    sat = opac.view(-1) > 1 - 1e-3
    assert int(sat.sum()) > 100
    # rays that did not saturate were marched to the end: same opacity as the one-shot render
    _mostly_close(opac[~sat], opac1[~sat], hard=0.5)
    assert float((opac[sat] <= opac1[sat] + 2e-3).float().mean()) > 0.99 and bool((opac <= 1 + 1e-5).all())
    assert np.isfinite(rgb.cpu().numpy()).all() and np.isfinite(depth.cpu().numpy()).all()


# ------------------------------------------------------------------ the fused loop (nerfacc_b200.chunked)
@pytest.mark.parametrize("cone,levels", [(0.0, 1), (0.004, 2)])
def test_fused_chunked_loop_chunks_are_the_reference_kernels_chunks(device, cone, levels):
    """nerfacc_b200.render_image_with_occgrid_test keeps the loop state on the device.  Every chunk
    it marches is compared, bit for bit, with the reference CUDA extension called the way the
    reference loop calls it (same alive mask, same near planes, same k): per-ray counts, sample
    order, t_starts / t_ends, termination planes; and k and the alive set follow utils.py:337,404-411."""
    import nerfacc_b200 as nf

    est, o, d = _scene(device, levels=levels)
    fn = _field(o, d)
    n = o.shape[0]
    step = 0.01
    far = torch.full((n,), 1e10, device=device)
    t_mins, t_maxs, hits = nf.ray_aabb_intersect(o, d, est.aabbs)
    t_sorted, t_indices = torch.sort(torch.cat([t_mins, t_maxs], -1), -1)
    min_samples = 1 if cone == 0 else 4
    seen = {"chunks": 0, "prev_alive": None}
    opacity_probe = {}

    def on_chunk(c):
        k, alive = c["k"], c["alive"].long()
        assert k == max(min(n // alive.numel(), 64), min_samples)
        mask = torch.zeros(n, dtype=torch.bool, device=device)
        mask[alive] = True
        assert int(mask.sum()) == alive.numel()  # no duplicates in the list
        pi = c["packed_info"]
        assert bool((pi[~mask, 1] == 0).all())
        if R.available():
            iv_r, sm_r, term_r = R.mod().traverse_grids(
                o, d, mask, est.binaries, est.aabbs, t_sorted, t_indices, hits, c["near_before"], far,
                float(step), float(cone), True, True, True, int(k), True)
            assert torch.equal(pi[:, 1], sm_r.chunk_cnts)
            assert torch.equal(pi[:, 0], sm_r.chunk_starts)
            assert torch.equal(c["t_starts"], iv_r.vals[iv_r.is_left])
            assert torch.equal(c["t_ends"], iv_r.vals[iv_r.is_right])
            assert torch.equal(c["ray_indices"], sm_r.ray_indices[sm_r.is_valid])
            assert torch.equal(c["near_after"][mask], term_r[mask])
            assert torch.equal(c["near_after"][~mask], c["near_before"][~mask])
        seen["chunks"] += 1

    rgb, opac, depth, total = nf.render_image_with_occgrid_test(
        4096, fn, est, o, d, render_step_size=step, cone_angle=cone, early_stop_eps=-1.0, _on_chunk=on_chunk)
    assert seen["chunks"] > 3
    with torch.no_grad():
        ri, ts, te = est.sampling(o, d, render_step_size=step, cone_angle=cone)
        colors1, opac1, depth1, _ = nf.rendering(ts, te, ri, n, rgb_sigma_fn=fn)
        # and the loop spelled with the public API (the reference's own sequence of calls)
        rgb2, opac2, depth2, total2, iters2 = _chunked(nf, est, o, d, fn, step, cone, early_stop_eps=-1.0)
    assert total == total2 and seen["chunks"] == iters2
    assert abs(total - ri.shape[0]) <= 1e-3 * ri.shape[0]
    _mostly_close(rgb, colors1)
    _mostly_close(opac, opac1)
    _mostly_close(depth, depth1, tol=5e-3, hard=0.3)
    # against the spelled-out loop: same chunks, so only summation order differs
    torch.testing.assert_close(rgb, rgb2, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(opac, opac2, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(depth, depth2 / opac2.clamp_min(torch.finfo(torch.float32).eps), rtol=1e-4, atol=1e-4)


def test_fused_chunked_loop_early_stop_alpha_thre_and_cost(device):
    import nerfacc_b200 as nf
    from nerfacc_b200 import _C

    est, o, d = _scene(device, seed=3)
    fn0 = _field(o, d)
    fn = lambda a, b, r: (fn0(a, b, r)[0], fn0(a, b, r)[1] * 40.0)  # dense: rays saturate early
    n = o.shape[0]
    chunks = []
    l0 = _C.launch_count()
    bk = torch.tensor([0.2, 0.4, 0.6], device=device)
    rgb, opac, depth, total = nf.render_image_with_occgrid_test(
        4096, fn, est, o, d, render_step_size=0.01, early_stop_eps=1e-3, alpha_thre=1e-2, render_bkgd=bk,
        _on_chunk=lambda c: chunks.append(c["k"]))
    launches = _C.launch_count() - l0
    # reference semantics spelled with the public API, including the alpha filter (utils.py:382-390)
    with torch.no_grad():
        opacity2 = torch.zeros((n, 1), device=device); depth2 = torch.zeros((n, 1), device=device)
        rgb2 = torch.zeros((n, 3), device=device)
        alive = torch.ones(n, dtype=torch.bool, device=device)
        near = torch.zeros(n, device=device); far = torch.full((n,), 1e10, device=device)
        iters = total2 = 0
        ks = []
        while iters < 4096:
            n_alive = int(alive.sum())
            if n_alive == 0:
                break
            k = max(min(n // n_alive, 64), 1)
            iters += k
            ks.append(k)
            intervals, samples, term = nf.traverse_grids(o, d, est.binaries, est.aabbs, near, far, 0.01, 0.0, k, True, alive)
            ts = intervals.vals[intervals.is_left]; te = intervals.vals[intervals.is_right]
            ri = samples.ray_indices[samples.is_valid]
            rgbs, sigmas = fn(ts, te, ri)
            w, _, a = nf.render_weight_from_density(ts, te, sigmas, ray_indices=ri, n_rays=n,
                                                    prefix_trans=1 - opacity2[ri].squeeze(-1))
            vis = a >= 1e-2
            nf.accumulate_along_rays_(w[vis], values=rgbs[vis], ray_indices=ri[vis], outputs=rgb2)
            nf.accumulate_along_rays_(w[vis], values=None, ray_indices=ri[vis], outputs=opacity2)
            nf.accumulate_along_rays_(w[vis], values=(ts[vis] + te[vis])[..., None] / 2.0, ray_indices=ri[vis], outputs=depth2)
            near = term
            alive = torch.logical_and(opacity2.view(-1) <= 1 - 1e-3, samples.packed_info[:, 1] == k)
            total2 += int(ri.shape[0])
        rgb2 = rgb2 + bk * (1.0 - opacity2)
        depth2 = depth2 / opacity2.clamp_min(torch.finfo(torch.float32).eps)
    # a ray sitting within rounding of the opacity threshold may retire one chunk apart in the two loops
    assert abs(total - total2) <= 2e-3 * total2 and abs(len(chunks) - len(ks)) <= 1
    same = (opac - opacity2).abs().view(-1) <= 1e-4
    assert float(same.float().mean()) > 0.999
    torch.testing.assert_close(rgb[same], rgb2[same], rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(depth[same], depth2[same], rtol=1e-3, atol=1e-3)
    assert int((opac.view(-1) > 1 - 1e-3).sum()) > 100 and bool((opac <= 1 + 1e-5).all())
    # cost: the fused loop issues at most 8 native launches per iteration (march, scan, compact,
    # composite, plan + first-call packing) and one host round trip; the spelled-out loop above makes
    # 5 syncs (.sum(), traverse totals, three mask selects) and ~25 launches per iteration
    assert launches <= 8 * len(chunks) + 8, (launches, len(chunks))
