"""Grid traversal on the GPU against the CPU oracle: ray_indices, packed_info, counts, masks and
edge values must be BIT-EXACT (same serial float arithmetic), as north_star demands.
Scenes follow the reference's tests/test_grid.py plus the cases it does not cover
(cone_angle > 0, step_size <= 0, zero direction components, chunked over-allocate mode)."""
import numpy as np
import pytest
import torch

import oracle

pytestmark = pytest.mark.gpu


def _enlarge(aabb, f):
    c, e = (aabb[:3] + aabb[3:]) / 2, (aabb[3:] - aabb[:3]) / 2
    return np.concatenate([c - e * f, c + e * f]).astype(np.float32)


def _nested_scene(n_rays=200, levels=4, res=32, seed=42, p_occ=0.5):
    rng = np.random.default_rng(seed)
    o = rng.standard_normal((n_rays, 3)).astype(np.float32)
    d = rng.standard_normal((n_rays, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=-1, keepdims=True)
    base = np.array([-1, -1, -1, 1, 1, 1], np.float32)
    aabbs = np.stack([_enlarge(base, 2**i) for i in range(levels)])
    bins = rng.random((levels, res, res, res)) < p_occ
    return o, d.astype(np.float32), bins, aabbs


def _shell_scene(n_rays=1 << 13, res=128, seed=1234):
    """thin spherical shell in a 128^3 grid, rays from radius 4 toward the ball."""
    rng = np.random.default_rng(seed)
    lin = (np.arange(res) + 0.5) / res * 3.0 - 1.5
    X, Y, Z = np.meshgrid(lin, lin, lin, indexing="ij")
    rad = np.sqrt(X**2 + Y**2 + Z**2)
    bins = ((rad >= 0.8) & (rad <= 0.9))[None]
    o = rng.standard_normal((n_rays, 3)); o = 4.0 * o / np.linalg.norm(o, axis=-1, keepdims=True)
    tgt = rng.standard_normal((n_rays, 3)); tgt = tgt / np.linalg.norm(tgt, axis=-1, keepdims=True)
    tgt *= 1.5 * rng.random((n_rays, 1)) ** (1 / 3)
    d = tgt - o; d /= np.linalg.norm(d, axis=-1, keepdims=True)
    aabbs = np.array([[-1.5, -1.5, -1.5, 1.5, 1.5, 1.5]], np.float32)
    return o.astype(np.float32), d.astype(np.float32), bins, aabbs


def _gpu(device, *arrs):
    return [None if a is None else torch.from_numpy(np.ascontiguousarray(a)).to(device) for a in arrs]


def _check_equal(got, ref):
    for part in ("intervals", "samples"):
        for k, v in ref[part].items():
            g = getattr(got[part], k).cpu().numpy()
            assert g.shape == v.shape, (part, k, g.shape, v.shape)
            assert (g == v).all(), (part, k, int((g != v).sum()))


@pytest.mark.parametrize("kw", [
    dict(),                                         # reference test_traverse_grids
    dict(step_size=0.013, cone_angle=0.004),        # cone marching
    dict(step_size=0.0),                            # one interval per voxel crossing
    dict(step_size=2e-2, traverse_steps_limit=37),  # per-ray sample cap, two-pass
])
def test_traverse_grids_matches_oracle(device, kw):
    import nerfacc_b200 as nf

    o, d, bins, aabbs = _nested_scene()
    near = np.random.default_rng(1).random(o.shape[0]).astype(np.float32) * 0.2
    far = np.full(o.shape[0], 6.0, np.float32)
    ref = oracle.traverse_grids(o, d, bins, aabbs, near, far, **kw)
    to, td, tb, ta, tn, tf = _gpu(device, o, d, bins, aabbs, near, far)
    iv, sm, term = nf.traverse_grids(to, td, tb, ta, near_planes=tn, far_planes=tf, **kw)
    _check_equal(dict(intervals=iv, samples=sm), ref)
    has = ref["samples"]["packed_info"][:, 1] > 0  # the reference leaves empty rays undefined
    assert (term.cpu().numpy()[has] == ref["termination_planes"][has]).all()
    assert ref["samples"]["vals"].shape[0] > 1000


def test_traverse_precomputed_events_and_zero_dirs(device):
    import nerfacc_b200 as nf

    o, d, bins, aabbs = _nested_scene(n_rays=64, levels=3)
    d[:8] = np.array([1.0, 0.0, 0.0], np.float32)
    d[8:16] = np.array([0.0, -1.0, 0.0], np.float32)
    d[16:20] = np.array([0.0, 0.6, 0.8], np.float32)
    o[:20] = (np.random.default_rng(3).random((20, 3)) * 1.6 - 0.8).astype(np.float32)
    ref = oracle.traverse_grids(o, d, bins, aabbs, step_size=5e-3)
    to, td, tb, ta = _gpu(device, o, d, bins, aabbs)
    iv, sm, _ = nf.traverse_grids(to, td, tb, ta, step_size=5e-3)
    _check_equal(dict(intervals=iv, samples=sm), ref)
    # same events handed in explicitly (the reference's call pattern at examples/utils.py:317-361)
    tmin, tmax, hits = nf.ray_aabb_intersect(to, td, ta)
    t_sorted, t_idx = torch.sort(torch.cat([tmin, tmax], -1), dim=-1, stable=True)
    iv2, sm2, _ = nf.traverse_grids(to, td, tb, ta, step_size=5e-3, t_sorted=t_sorted, t_indices=t_idx, hits=hits)
    _check_equal(dict(intervals=iv2, samples=sm2), ref)


def test_ray_aabb_intersect(device):
    """Reference tests/test_grid.py:8-35 + bit-exactness against the oracle."""
    import nerfacc_b200 as nf
    from nerfacc_b200.grid import _ray_aabb_intersect

    rng = np.random.default_rng(42)
    o = rng.random((1000, 3)).astype(np.float32)
    d = rng.standard_normal((1000, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=-1, keepdims=True)
    lo = rng.random((100, 3)).astype(np.float32)
    bb = np.concatenate([lo, lo + rng.random((100, 3)).astype(np.float32)], -1)
    to, td, tb = _gpu(device, o, d, bb)
    tmin, tmax, hits = nf.ray_aabb_intersect(to, td, tb)
    r0, r1, rh = oracle.ray_aabb_intersect(o, d, bb)
    assert (hits.cpu().numpy() == rh).all()
    assert (tmin.cpu().numpy() == r0).all() and (tmax.cpu().numpy() == r1).all()
    _a, _b, _h = _ray_aabb_intersect(to, td, tb)
    assert torch.allclose(tmin, _a) and torch.allclose(tmax, _b) and (hits == _h).all()
    tmin, tmax, hits = nf.ray_aabb_intersect(to, td, tb, near_plane=0.1, far_plane=0.7, miss_value=-1.0)
    r0, r1, rh = oracle.ray_aabb_intersect(o, d, bb, 0.1, 0.7, -1.0)
    assert (tmin.cpu().numpy() == r0).all() and (tmax.cpu().numpy() == r1).all()


def test_near_far_planes_kat(device):
    """Reference tests/test_grid.py:135-159 (1x1x1 grid)."""
    import nerfacc_b200 as nf

    rays_o = torch.tensor([[-1.0, 0.0, 0.0]], device=device)
    rays_d = torch.tensor([[1.0, 0.01, 0.01]], device=device)
    rays_d = rays_d / rays_d.norm(dim=-1, keepdim=True)
    binaries = torch.ones((1, 1, 1, 1), dtype=torch.bool, device=device)
    aabbs = torch.tensor([[0.0, 0.0, 0.0, 1.0, 1.0, 1.0]], device=device)
    near, far = torch.tensor([1.2], device=device), torch.tensor([1.5], device=device)
    iv, sm, _ = nf.traverse_grids(rays_o, rays_d, binaries, aabbs, step_size=0.05, near_planes=near, far_planes=far)
    assert iv.vals.numel() > 0
    assert (iv.vals >= (near - 0.025)).all() and (iv.vals <= (far + 0.025)).all()
    ref = oracle.traverse_grids(rays_o.cpu().numpy(), rays_d.cpu().numpy(), binaries.cpu().numpy(),
                                aabbs.cpu().numpy(), near.cpu().numpy(), far.cpu().numpy(), step_size=0.05)
    assert (iv.vals.cpu().numpy() == ref["intervals"]["vals"]).all()


def test_chunked_test_mode(device):
    """Reference tests/test_grid.py:72-131: over-allocated, resumable marching reproduces the
    one-shot result; every chunk is also bit-exact against the oracle."""
    import nerfacc_b200 as nf

    o, d, bins, aabbs = _nested_scene(n_rays=10)
    to, td, tb, ta = _gpu(device, o, d, bins, aabbs)
    iv, sm, _ = nf.traverse_grids(to, td, tb, ta)
    one_shot = nf.accumulate_along_rays(iv.vals[iv.is_left], None, sm.ray_indices, 10)
    acc = torch.zeros_like(one_shot)
    term, mask, term_np, mask_np = None, None, None, None
    for _ in range(3):
        _iv, _sm, term = nf.traverse_grids(to, td, tb, ta, near_planes=term, traverse_steps_limit=4000,
                                           over_allocate=True, rays_mask=mask)
        ref = oracle.traverse_grids(o, d, bins, aabbs, near_planes=term_np, traverse_steps_limit=4000,
                                    over_allocate=True, rays_mask=mask_np)
        _check_equal(dict(intervals=_iv, samples=_sm), ref)
        alive = ref["samples"]["packed_info"][:, 1] > 0
        assert (term.cpu().numpy()[alive] == ref["termination_planes"][alive]).all()
        mask = _sm.packed_info[:, 1] == 4000
        mask_np, term_np = mask.cpu().numpy(), term.cpu().numpy()
        nf.accumulate_along_rays_(_iv.vals[_iv.is_left], None, _sm.ray_indices[_sm.is_valid], acc)
    assert (~mask).all()
    assert torch.allclose(acc, one_shot, atol=1e-1)


@pytest.mark.parametrize("stratified", [False, True])
def test_sampling_shell_scene(device, stratified):
    """cfg2-shaped scene at 2^13 rays: ray_indices / t_starts / t_ends / counts bit-exact."""
    import nerfacc_b200 as nf

    o, d, bins, aabbs = _shell_scene()
    est = nf.OccGridEstimator(aabbs[0].tolist(), resolution=128, levels=1).to(device)
    est.binaries = torch.from_numpy(bins).to(device)
    to, td = _gpu(device, o, d)
    if stratified:
        torch.manual_seed(7)
        jitter = torch.rand(o.shape[0], device=device)
        torch.manual_seed(7)
    ri, ts, te = est.sampling(to, td, render_step_size=5e-3, stratified=stratified)
    near = None
    if stratified:
        near = (torch.zeros(o.shape[0], device=device) + jitter * 5e-3).cpu().numpy()
    r_ri, r_ts, r_te, r_pi = oracle.sampling(o, d, bins, aabbs, render_step_size=5e-3, t_min=near)
    assert ri.shape[0] == r_ri.shape[0] and ri.shape[0] > 100000
    assert (ri.cpu().numpy() == r_ri).all()
    assert (ts.cpu().numpy() == r_ts).all() and (te.cpu().numpy() == r_te).all()
    assert (nf.pack_info(ri, o.shape[0]).cpu().numpy() == r_pi).all()


def test_sampling_with_min_max_and_visibility(device):
    """Reference tests/test_grid.py:163-203 plus the sigma_fn visibility filter."""
    import nerfacc_b200 as nf

    rng = np.random.default_rng(42)
    n_rays, levels, res = 64, 4, 32
    o = (rng.random((n_rays, 3)) * 2 - 1).astype(np.float32)
    d = rng.random((n_rays, 3)).astype(np.float32); d /= np.linalg.norm(d, axis=-1, keepdims=True)
    bins = rng.random((levels, res, res, res)) > 0.5
    t_min = rng.random(n_rays).astype(np.float32)
    t_max = (t_min + rng.random(n_rays)).astype(np.float32)
    est = nf.OccGridEstimator([-1.0, -1.0, -1.0, 1.0, 1.0, 1.0], resolution=res, levels=levels).to(device)
    est.binaries = torch.from_numpy(bins).to(device)
    to, td, tmn, tmx = _gpu(device, o, d, t_min, t_max)
    ri, ts, te = est.sampling(to, td, near_plane=0.15, far_plane=0.85, t_min=tmn, t_max=tmx, render_step_size=0.01)
    assert (ts >= (tmn[ri] - 0.005)).all() and (te <= (tmx[ri] + 0.005)).all()
    aab = est.aabbs.cpu().numpy()
    r_ri, r_ts, r_te, r_pi = oracle.sampling(o, d, bins, aab, 0.15, 0.85, t_min, t_max, 0.01)
    assert (ri.cpu().numpy() == r_ri).all() and (ts.cpu().numpy() == r_ts).all() and (te.cpu().numpy() == r_te).all()
    # visibility filter: sigma = 30 everywhere -> compare with oracle weights + mask
    est.occs.fill_(1.0)
    ri2, ts2, te2 = est.sampling(to, td, sigma_fn=lambda a, b, r: torch.full_like(a, 30.0), near_plane=0.15,
                                 far_plane=0.85, t_min=tmn, t_max=tmx, render_step_size=0.01,
                                 early_stop_eps=1e-2, alpha_thre=1e-3)
    w0, T0, a0 = oracle.render_weight_from_density(r_ts, r_te, np.full_like(r_ts, 30.0), r_pi)
    keep = oracle.visibility(T0, a0, 1e-2, 1e-3)
    safe = np.abs(T0 - 1e-2) > 1e-6
    assert safe.all()
    assert (ri2.cpu().numpy() == r_ri[keep]).all() and (ts2.cpu().numpy() == r_ts[keep]).all()
    assert (te2.cpu().numpy() == r_te[keep]).all()


def test_mark_invisible_cells(device):
    """Reference tests/test_grid.py:207-233 (exact counts)."""
    import nerfacc_b200 as nf

    est = nf.OccGridEstimator([-1.0, -1.0, -1.0, 1.0, 1.0, 1.0], resolution=32, levels=4).to(device)
    K = torch.tensor([[[100.0, 0, 50.0], [0, 100.0, 50.0], [0, 0, 1]]], device=device)
    pose = torch.tensor([[[-1.0, 0.0, 0.0, 0.0], [0.0, 1.0, 0.0, 0.0], [0.0, 0.0, -1.0, 2.5]]], device=device)
    est.mark_invisible_cells(K, pose, 100, 100)
    assert (est.occs == -1).sum() == 77660
    assert (est.occs == 0).sum() == 53412


def test_update_every_n_steps(device):
    """EMA update + binarisation against the oracle, with an occ_eval_fn that is constant per
    cell (so the reference's last-writer-wins and our scatter-max agree)."""
    import nerfacc_b200 as nf

    res, levels = 32, 2
    est = nf.OccGridEstimator([-1.0, -1.0, -1.0, 1.0, 1.0, 1.0], resolution=res, levels=levels).to(device)
    est.train()
    aab = est.aabbs.cpu().numpy()

    def occ_fn(x):  # depends only on which level-0-sized cell x is in
        cell = torch.floor((x + 2.0) / 4.0 * 64).clamp(0, 63)
        h = (cell[:, 0] * 7 + cell[:, 1] * 13 + cell[:, 2] * 29) % 17
        return (h[:, None] / 17.0 * 0.05).float()

    occs = np.zeros(levels * res**3, np.float32)
    for step in (0, 16, 512, 528):  # warm-up and steady-state sampling both run
        est.update_every_n_steps(step, occ_fn, occ_thre=0.01, ema_decay=0.95, warmup_steps=256)
    # warm-up steps touch every cell once: closed form check against the oracle
    est2 = nf.OccGridEstimator([-1.0, -1.0, -1.0, 1.0, 1.0, 1.0], resolution=res, levels=levels).to(device)
    est2.train()
    torch.manual_seed(5)
    est2._update(0, occ_fn, occ_thre=0.01, ema_decay=0.95, warmup_steps=256)
    torch.manual_seed(5)
    for lvl in range(levels):
        cells = np.arange(res**3)
        u = torch.rand((res**3, 3), device=device).cpu().numpy()
        x = oracle.occ_cell_points(cells, u, [res] * 3, aab[lvl])
        occ = occ_fn(torch.from_numpy(x).to(device)).squeeze(-1).cpu().numpy()
        occs = oracle.occ_ema_update(occs, lvl * res**3 + cells, occ, 0.95)
    bins, thre = oracle.occ_binarize(occs, 0.01)
    assert np.allclose(est2.occs.cpu().numpy(), occs, rtol=0, atol=0)
    assert (est2.binaries.cpu().numpy().reshape(-1) == bins).all()
    assert est.binaries.any() and not est.binaries.all()


def test_ema_update_with_duplicates_and_compaction(device):
    from nerfacc_b200 import _occ_ops

    rng = np.random.default_rng(0)
    n_cells, n = 5000, 20000
    occs = rng.random(n_cells).astype(np.float32)
    occs[rng.random(n_cells) < 0.1] = -1.0
    ids = rng.integers(0, n_cells, n)
    occ = rng.random(n).astype(np.float32)
    t_occs = torch.from_numpy(occs.copy()).to(device)
    _occ_ops.ema_update(t_occs, torch.from_numpy(ids).to(device), torch.from_numpy(occ).to(device), 0, n, 0.9)
    ref = oracle.occ_ema_update(occs, ids, occ, 0.9)
    assert (t_occs.cpu().numpy() == ref).all()
    b = _occ_ops.binarize(t_occs, 0.5).cpu().numpy()
    rb, _ = oracle.occ_binarize(ref, 0.5)
    assert (b == rb).all()
    # compaction
    m = 100003
    mask = rng.random(m) < 0.3
    ri = np.sort(rng.integers(0, 999, m)); ts = rng.random(m).astype(np.float32); te = ts + 1
    a, b2, c = _occ_ops.compact_samples(*[torch.from_numpy(x).to(device) for x in (mask, ri, ts, te)], 999)
    assert (a.cpu().numpy() == ri[mask]).all() and (b2.cpu().numpy() == ts[mask]).all() and (c.cpu().numpy() == te[mask]).all()


def test_sampling_run_pool_overflow(device):
    """Busy grid (40 % random cells): dozens of runs per ray, far beyond the 8 inline slots and
    the run pool (2 per ray), so most rays take the re-march work list — bit-exact all the same."""
    import nerfacc_b200 as nf

    rng = np.random.default_rng(123)
    n_rays, res = 5000, 48
    o = (rng.standard_normal((n_rays, 3)) * 0.3).astype(np.float32)
    d = rng.standard_normal((n_rays, 3)).astype(np.float32); d /= np.linalg.norm(d, axis=-1, keepdims=True)
    bins = rng.random((1, res, res, res)) < 0.4
    est = nf.OccGridEstimator([-1.0, -1.0, -1.0, 1.0, 1.0, 1.0], resolution=res, levels=1).to(device)
    est.binaries = torch.from_numpy(bins).to(device)
    to, td = _gpu(device, o, d)
    ri, ts, te = est.sampling(to, td, render_step_size=4e-3)
    r_ri, r_ts, r_te, r_pi = oracle.sampling(o, d, bins, est.aabbs.cpu().numpy(), render_step_size=4e-3)
    assert ri.shape[0] == r_ri.shape[0] > 50 * n_rays
    assert (ri.cpu().numpy() == r_ri).all()
    assert (ts.cpu().numpy() == r_ts).all() and (te.cpu().numpy() == r_te).all()


def test_occupancy_bounds(device):
    """nfb_occupancy_bounds vs numpy: levels whose cells do not fill whole 32-bit words or whole
    warps of words, an empty level, a single cell."""
    from nerfacc_b200 import _grid_ops

    rng = np.random.default_rng(3)
    for res in [(5, 7, 9), (33, 31, 37), (128, 128, 128)]:
        bins = np.zeros((4,) + res, bool)
        bins[0] = rng.random(res) < 0.01
        bins[2, res[0] // 2, res[1] - 1, 0] = True
        lo, hi = [r // 4 for r in res], [r // 2 + 1 for r in res]
        bins[3, lo[0]:hi[0], lo[1]:hi[1], lo[2]:hi[2]] = rng.random([h - l for l, h in zip(lo, hi)]) < 0.3
        _, _, _, bounds = _grid_ops.occupancy_bits(torch.from_numpy(bins).to(device))
        got = bounds.cpu().numpy()
        for l in range(4):
            idx = np.argwhere(bins[l])
            want = list(res) + [-1, -1, -1] if len(idx) == 0 else list(idx.min(0)) + list(idx.max(0))
            assert got[l].tolist() == want, (res, l, got[l], want)


@pytest.mark.parametrize("levels", [1, 3])
def test_sampling_stops_at_occupied_box(device, levels):
    """Marches that stop walking at the box of the occupied cells (grid.cuh setup_dda `clip`):
    small off-centre blobs, an empty outer level, rays starting inside and outside the boxes and
    moving every way — samples bit-exact against the oracle's full walk."""
    import nerfacc_b200 as nf

    res, n_rays = 32, 6000
    rng = np.random.default_rng(17)
    bins = np.zeros((levels, res, res, res), bool)
    bins[0, 20:27, 3:9, 11:14] = rng.random((7, 6, 3)) < 0.6
    bins[0, 5, 25, 30] = True
    if levels > 1:
        bins[1, 2:6, 10:30, 28:32] = rng.random((4, 20, 4)) < 0.5  # level 2 stays empty
    base = np.array([-1, -1, -1, 1, 1, 1], np.float32)
    aabbs = np.stack([_enlarge(base, 2**i) for i in range(levels)])
    o = (rng.random((n_rays, 3)) * 6 - 3).astype(np.float32)
    o[: n_rays // 3] = (rng.random((n_rays // 3, 3)) * 1.6 - 0.8).astype(np.float32)  # inside level 0
    d = rng.standard_normal((n_rays, 3)).astype(np.float32)
    d[::7, 0] = 0.0
    d[::11, 2] = 0.0
    d /= np.linalg.norm(d, axis=-1, keepdims=True)
    est = nf.OccGridEstimator(base.tolist(), resolution=res, levels=levels).to(device)
    est.binaries = torch.from_numpy(bins).to(device)
    to, td = _gpu(device, o, d)
    ri, ts, te = est.sampling(to, td, render_step_size=7e-3)
    r_ri, r_ts, r_te, r_pi = oracle.sampling(o, d, bins, aabbs, render_step_size=7e-3)
    assert ri.shape[0] == r_ri.shape[0] and ri.shape[0] > 1000
    assert (ri.cpu().numpy() == r_ri).all()
    assert (ts.cpu().numpy() == r_ts).all() and (te.cpu().numpy() == r_te).all()


@pytest.mark.parametrize("near,step", [(0.0, 5e-3), (0.05, 0.0123), (1.3, 1.5 * 2.0**-8), (0.0, 2.0**-7),
                                       (2.9, 3e-3), (0.0, 0.25)])
def test_sampling_long_skips_from_checkpoints(device, near, step):
    """Constant dt from one near plane: long skips start from a checkpoint of the common t
    sequence (grid.cuh make_checkpoints) — bit-exact against the oracle's step-by-step walk for
    near planes before / inside the scene and step sizes with short and long mantissas."""
    import nerfacc_b200 as nf

    o, d, bins, aabbs = _shell_scene(n_rays=1 << 12, res=64, seed=21)
    est = nf.OccGridEstimator(aabbs[0].tolist(), resolution=64, levels=1).to(device)
    est.binaries = torch.from_numpy(bins).to(device)
    to, td = _gpu(device, o, d)
    ri, ts, te = est.sampling(to, td, render_step_size=step, near_plane=near)
    r_ri, r_ts, r_te, r_pi = oracle.sampling(o, d, bins, aabbs, near_plane=near, render_step_size=step)
    assert ri.shape[0] == r_ri.shape[0] and ri.shape[0] > 100
    assert (ri.cpu().numpy() == r_ri).all()
    assert (ts.cpu().numpy() == r_ts).all() and (te.cpu().numpy() == r_te).all()


def test_update_steady_state_vs_oracle(device):
    """The steady-state branch of _update (step >= warmup_steps: res^3/4 uniform draws with
    replacement — duplicates — plus the occupied cells, occ_grid.py:348-365, 367-404) through the
    estimator, compared cell for cell with the oracle fed the same draws; and the reference's own
    OccGridEstimator._update (pure torch, last-writer-wins) on the same RNG stream where the
    reference package is present — with an occ_eval_fn that is constant per cell the two semantics
    coincide."""
    import nerfacc_b200 as nf

    res, levels = 32, 2
    roi = [-1.0, -1.0, -1.0, 1.0, 1.0, 1.0]

    def occ_fn(x):  # depends only on which level-0-sized cell x is in
        cell = torch.floor((x + 2.0) / 4.0 * 64).clamp(0, 63)
        h = (cell[:, 0] * 7 + cell[:, 1] * 13 + cell[:, 2] * 29) % 17
        return (h[:, None] / 17.0 * 0.05).float()

    est = nf.OccGridEstimator(roi, resolution=res, levels=levels).to(device)
    est.train()
    est.occs[::11] = -1.0  # invisible cells: never sampled, stay -1
    torch.manual_seed(3)
    est._update(0, occ_fn)  # a warm-up update makes some cells occupied
    assert est.binaries.any()
    for step, seed in ((256, 7), (272, 8)):
        occs0 = est.occs.cpu().numpy().copy()
        bins0 = est.binaries.clone()
        torch.manual_seed(seed)
        est._update(step, occ_fn, occ_thre=0.01, ema_decay=0.95, warmup_steps=256)
        # replay the draws: same generator state, same order of calls as _update makes
        torch.manual_seed(seed)
        shadow = nf.OccGridEstimator(roi, resolution=res, levels=levels).to(device)
        shadow.occs = torch.from_numpy(occs0).to(device)
        shadow.binaries = bins0
        lvl_cells = shadow._sample_uniform_and_occupied_cells(res**3 // 4)
        occs = occs0.copy()
        n_dup = 0
        for lvl, cells in enumerate(lvl_cells):
            u = torch.rand((cells.shape[0], 3), device=device)
            c = cells.cpu().numpy()
            n_dup += c.size - np.unique(c).size
            assert (occs0[lvl * res**3 + c] >= 0).all()
            x = oracle.occ_cell_points(c, u.cpu().numpy(), [res] * 3, est.aabbs[lvl].cpu().numpy())
            occ = occ_fn(torch.from_numpy(x).to(device)).squeeze(-1).cpu().numpy()
            occs = oracle.occ_ema_update(occs, lvl * res**3 + c, occ, 0.95)
        assert n_dup > 100  # the draw is with replacement
        bins, _ = oracle.occ_binarize(occs, 0.01)
        assert np.array_equal(est.occs.cpu().numpy(), occs)
        assert np.array_equal(est.binaries.cpu().numpy().reshape(-1), bins.astype(bool))
        assert (est.occs[::11] == -1.0).all()


@pytest.mark.parametrize("n_cells,n,frac_occ", [(32**3, 32**3 // 4, 0.05), (20 * 21 * 23, 777, 0.6),
                                                (5000, 0, 0.3), (4099, 9000, 0.0), (2048, 2048, 1.0)])
def test_select_cells_vs_torch(device, n_cells, n, frac_occ):
    """nfb_occ_select_cells (two launches, one read-back) against the reference's formulation
    (occ_grid.py:344-365): uniform[occs[uniform] >= 0] followed by nonzero(binaries), in order;
    sizes that are not multiples of the tile, no draws, no occupied cells, everything occupied."""
    from nerfacc_b200 import _occ_ops

    g = torch.Generator(device=device).manual_seed(n_cells + n)
    occs = torch.rand(n_cells, device=device, generator=g)
    occs[torch.rand(n_cells, device=device, generator=g) < 0.2] = -1.0
    bins = torch.rand(n_cells, device=device, generator=g) < frac_occ
    uniform = torch.randint(n_cells, (n,), device=device, generator=g)
    cells, k1, k2 = _occ_ops.select_cells(occs, uniform, bins)
    want_u = uniform[occs[uniform] >= 0.0]
    want_o = torch.nonzero(bins)[:, 0]
    assert (k1, k2) == (want_u.numel(), want_o.numel())
    assert torch.equal(cells[:k1], want_u) and torch.equal(cells[k1:k1 + k2], want_o)


def test_sample_cells_more_occupied_than_draws(device):
    """The branch of _sample_uniform_and_occupied_cells that subsamples the occupied cells
    (n < len(occupied), occ_grid.py:358-362): same RNG stream, same cells as the torch formulation."""
    import nerfacc_b200 as nf

    res, levels = 16, 2
    est = nf.OccGridEstimator([-1.0] * 3 + [1.0] * 3, resolution=res, levels=levels).to(device)
    g = torch.Generator(device=device).manual_seed(5)
    est.binaries = torch.rand((levels, res, res, res), device=device, generator=g) < 0.7
    est.occs = torch.where(torch.rand(levels * res**3, device=device, generator=g) < 0.3, -1.0, 0.5)
    n = res**3 // 4
    torch.manual_seed(11)
    got = est._sample_uniform_and_occupied_cells(n)
    torch.manual_seed(11)
    occs = est.occs.view(levels, -1)
    for lvl in range(levels):
        uniform = torch.randint(res**3, (n,), device=device)
        uniform = uniform[occs[lvl][uniform] >= 0.0]
        occupied = torch.nonzero(est.binaries[lvl].flatten())[:, 0]
        assert n < len(occupied)
        occupied = occupied[torch.randint(len(occupied), (n,), device=device)]
        assert torch.equal(got[lvl], torch.cat([uniform, occupied]))


def test_update_vs_reference_package(device):
    """OccGridEstimator._update, warm-up and steady state, against the reference's own estimator
    (baseline/_ref, pure torch on the same CUDA RNG stream): identical `occs` and `binaries`."""
    from tests import _refpkg

    if not _refpkg.available():
        pytest.skip("baseline/_ref (reference package) or oracle/_ref not present")
    import nerfacc_b200 as nf

    ref = _refpkg.load()
    res, levels = 32, 2
    roi = [-1.0, -1.0, -1.0, 1.0, 1.0, 1.0]

    def occ_fn(x):
        # constant inside every cell of BOTH levels (a level-1 cell is 1/8 wide: 32 of them across
        # [-2, 2]; level-0 cells nest inside): duplicates of a cell then all carry the same value and
        # the reference's last-writer-wins index_put equals the scatter-max
        cell = torch.floor((x + 2.0) / 4.0 * 32).clamp(0, 31)
        h = (cell[:, 0] * 7 + cell[:, 1] * 13 + cell[:, 2] * 29) % 17
        return (h[:, None] / 17.0 * 0.05).float()

    ours = nf.OccGridEstimator(roi, resolution=res, levels=levels).to(device)
    theirs = ref.OccGridEstimator(roi, resolution=res, levels=levels).to(device)
    ours.train(), theirs.train()
    for step in (0, 16, 256, 272, 288):
        torch.manual_seed(100 + step)
        ours.update_every_n_steps(step, occ_fn, occ_thre=0.01, ema_decay=0.95, warmup_steps=256)
        torch.manual_seed(100 + step)
        theirs.update_every_n_steps(step, occ_fn, occ_thre=0.01, ema_decay=0.95, warmup_steps=256)
        assert torch.equal(ours.binaries, theirs.binaries), step
        # occs: decay and max are the same float ops; the mean is only used for the threshold
        assert torch.equal(ours.occs, theirs.occs), step


def test_mark_invisible_cells_many_cameras_vs_reference_package(device):
    """The one-launch kernel against the reference estimator's own mark_invisible_cells
    (baseline/_ref; torch matmuls) with 300 random cameras, a near plane and a non-cubic grid.
    The per-cell decision is a chain of fp32 compares on projected coordinates; cuBLAS may
    accumulate the 3-term dot products in another order than the kernel's FMA chain, so a cell
    sitting within an ulp of an image border may flip: at most 1 in 10^4 cells may differ."""
    from tests import _refpkg

    if not _refpkg.available():
        pytest.skip("baseline/_ref (reference package) or oracle/_ref not present")
    import nerfacc_b200 as nf

    ref = _refpkg.load()
    g = torch.Generator(device=device).manual_seed(9)
    n_cams = 300
    eye = torch.randn((n_cams, 3), device=device, generator=g)
    eye = 2.5 * eye / eye.norm(dim=-1, keepdim=True)
    z = -eye / eye.norm(dim=-1, keepdim=True)  # look at the origin
    up = torch.tensor([0.0, 0.0, 1.0], device=device).expand_as(z)
    x = torch.cross(up, z, dim=-1)
    x = x / x.norm(dim=-1, keepdim=True)
    y = torch.cross(z, x, dim=-1)
    c2w = torch.cat([torch.stack([x, y, z], dim=-1), eye[:, :, None]], dim=-1)  # (n, 3, 4)
    K = torch.tensor([[[60.0, 0, 32.0], [0, 60.0, 24.0], [0, 0, 1]]], device=device).repeat(n_cams, 1, 1)
    K[:, 0, 0] += torch.rand(n_cams, device=device, generator=g) * 20
    kw = dict(roi_aabb=[-1.0, -0.8, -0.6, 1.0, 0.8, 0.6], resolution=[40, 32, 24], levels=3)
    ours, theirs = nf.OccGridEstimator(**kw).to(device), ref.OccGridEstimator(**kw).to(device)
    ours.mark_invisible_cells(K, c2w, 64, 48, near_plane=1.2)
    theirs.mark_invisible_cells(K, c2w, 64, 48, near_plane=1.2)
    diff = int((ours.occs != theirs.occs).sum())
    assert diff <= ours.occs.numel() // 10000, diff
    assert 0 < int((ours.occs == -1).sum()) < ours.occs.numel()
    # a second call only looks at cells that are still visible (occ_grid.py:294 via _get_all_cells)
    before = ours.occs.clone()
    ours.mark_invisible_cells(K[:1], c2w[:1], 64, 48, near_plane=0.0)
    assert bool((ours.occs[before == -1] == -1).all())


def test_sampling_expands_ahead_of_the_count_readback(device):
    """From the second call on, sampling() expands the runs into buffers sized from the previous
    call's total before this call's total has reached the host (nfb_traverse_expand).  Results must
    be the same as the first (exact-size) call; a total that outgrows the guess, a shrinking one, and
    rays whose runs overflow the run buffer all take their fallbacks."""
    import nerfacc_b200 as nf

    rng = np.random.default_rng(7)
    n_rays, res = 20000, 64
    o = rng.standard_normal((n_rays, 3)); o = (3.0 * o / np.linalg.norm(o, axis=-1, keepdims=True)).astype(np.float32)
    d = -o + 0.4 * rng.standard_normal((n_rays, 3)); d = (d / np.linalg.norm(d, axis=-1, keepdims=True)).astype(np.float32)
    to, td = torch.from_numpy(o).to(device), torch.from_numpy(d).to(device)
    aabb = np.array([[-1, -1, -1, 1, 1, 1]], np.float32)
    sparse = rng.random((1, res, res, res)) < 0.02
    dense = rng.random((1, res, res, res)) < 0.45   # dozens of runs per ray: run-buffer overflow
    est = nf.OccGridEstimator(aabb[0].tolist(), resolution=res).to(device)

    def check(bins, step):
        est.binaries = torch.from_numpy(bins).to(device)
        ri, ts, te = est.sampling(to, td, render_step_size=step)
        r_ri, r_ts, r_te, r_pi = oracle.sampling(o, d, bins, aabb, render_step_size=step)
        assert ri.shape[0] == r_ri.shape[0]
        assert (ri.cpu().numpy() == r_ri).all() and (ts.cpu().numpy() == r_ts).all() and (te.cpu().numpy() == r_te).all()
        assert (nf.pack_info(ri, n_rays).cpu().numpy() == r_pi).all()
        return ri.shape[0]

    a = check(sparse, 0.01)       # first call: no guess yet
    assert est._count_hint[n_rays] == a
    assert check(sparse, 0.01) == a   # guess fits: speculative buffers kept
    b = check(dense, 0.01)        # total far above the guess: exact re-expansion, overflowed rays re-marched
    assert b > 4 * a
    assert check(dense, 0.01) == b    # guess fits, rays overflow the run buffer: only those are marched again
    assert check(sparse, 0.02) < a    # shrinking total


def test_cone_sampling_from_the_piece_log(device):
    """Cone stepping: pass 1 logs every run of samples as pieces of <= 32 samples and pass 2 replays
    them (nfb_traverse_expand_cone) instead of marching again.  Bit-exact against the oracle for: a
    shell scene (runs of hundreds of samples at small t, cut into many pieces), nested levels, the
    speculative expansion from the second call on, a total that outgrows the guess, and a busy grid
    whose pieces fill the log (more than 32 per ray) so that pass 2 falls back to marching."""
    import nerfacc_b200 as nf
    from nerfacc_b200 import _C

    rng = np.random.default_rng(11)
    n_rays, res = 6000, 64
    o = rng.standard_normal((n_rays, 3)); o = (2.5 * o / np.linalg.norm(o, axis=-1, keepdims=True)).astype(np.float32)
    d = -o + 0.5 * rng.standard_normal((n_rays, 3)); d = (d / np.linalg.norm(d, axis=-1, keepdims=True)).astype(np.float32)
    to, td = torch.from_numpy(o).to(device), torch.from_numpy(d).to(device)
    lin = (np.arange(res) + 0.5) / res * 2 - 1
    X, Y, Z = np.meshgrid(lin, lin, lin, indexing="ij")
    rad = np.sqrt(X * X + Y * Y + Z * Z)
    shell = ((rad > 0.6) & (rad < 0.9))
    kw = dict(render_step_size=1e-3, cone_angle=0.004, near_plane=0.05)

    def check(est, bins, aabbs, expect_calls):
        est.binaries = torch.from_numpy(bins).to(device)
        _C.prof_enable(True)
        try:
            ri, ts, te = est.sampling(to, td, **kw)
            torch.cuda.synchronize()
            names = list(_C.prof_collect().keys())  # in order of first call
        finally:
            _C.prof_enable(False)
        r_ri, r_ts, r_te, r_pi = oracle.sampling(o, d, bins, aabbs, **kw)
        assert ri.shape[0] == r_ri.shape[0] > 0
        assert (ri.cpu().numpy() == r_ri).all() and (ts.cpu().numpy() == r_ts).all() and (te.cpu().numpy() == r_te).all()
        assert [n for n in names if n.startswith("traverse")] == expect_calls, names
        return ri.shape[0]

    # one level
    est = nf.OccGridEstimator([-1.0, -1.0, -1.0, 1.0, 1.0, 1.0], resolution=res, levels=1).to(device)
    aabbs = est.aabbs.cpu().numpy()
    sparse = shell[None] & (rng.random((1, res, res, res)) < 0.5)
    a = check(est, shell[None], aabbs, ["traverse_count", "traverse_fill_sampling"])  # no guess yet: exact buffers
    assert a > 50 * n_rays  # long runs: cut into several pieces each
    assert check(est, shell[None], aabbs, ["traverse_count", "traverse_expand_cone"]) == a  # the guess fits
    b = check(est, sparse, aabbs, ["traverse_count", "traverse_expand_cone"])  # fewer samples, many short runs
    assert b < a
    # the total outgrows the guess: expanded ahead into too small buffers, then again into exact ones
    assert check(est, shell[None], aabbs, ["traverse_count", "traverse_expand_cone", "traverse_fill_sampling"]) == a
    # per-ray near / far planes (t_min / t_max): pieces start wherever the ray does
    t_min = (rng.random(n_rays) * 2.0).astype(np.float32)
    t_max = (t_min + 0.2 + rng.random(n_rays) * 2.0).astype(np.float32)
    est.binaries = torch.from_numpy(shell[None]).to(device)
    ri, ts, te = est.sampling(to, td, t_min=torch.from_numpy(t_min).to(device), t_max=torch.from_numpy(t_max).to(device), **kw)
    r_ri, r_ts, r_te, _ = oracle.sampling(o, d, shell[None], aabbs, t_min=t_min, t_max=t_max, **kw)
    assert ri.shape[0] == r_ri.shape[0] > 0
    assert (ri.cpu().numpy() == r_ri).all() and (ts.cpu().numpy() == r_ts).all() and (te.cpu().numpy() == r_te).all()
    # nested levels
    est3 = nf.OccGridEstimator([-0.5, -0.5, -0.5, 0.5, 0.5, 0.5], resolution=res, levels=3).to(device)
    bins3 = np.stack([shell | (rng.random((res, res, res)) < 0.01) for _ in range(3)])
    check(est3, bins3, est3.aabbs.cpu().numpy(), ["traverse_count", "traverse_fill_sampling"])
    check(est3, bins3, est3.aabbs.cpu().numpy(), ["traverse_count", "traverse_expand_cone"])
    # a grid so busy that the log runs full (runs of one or two cells, more than 32 pieces per ray):
    # pass 2 marches (the speculative expansion's output is discarded)
    est_b = nf.OccGridEstimator([-1.0, -1.0, -1.0, 1.0, 1.0, 1.0], resolution=128, levels=1).to(device)
    busy128 = rng.random((1, 128, 128, 128)) < 0.5
    n1 = check(est_b, busy128, est_b.aabbs.cpu().numpy(), ["traverse_count", "traverse_fill_sampling"])
    assert check(est_b, busy128, est_b.aabbs.cpu().numpy(),
                 ["traverse_count", "traverse_expand_cone", "traverse_fill_sampling"]) == n1


@pytest.mark.parametrize("n_rays", [0, 1, 33, 257])
@pytest.mark.parametrize("fill", ["empty", "full", "shell"])
def test_cone_sampling_small_and_degenerate(device, n_rays, fill):
    """The piece log at the edges: no rays, fewer rays than a warp, a grid without any occupied cell
    (no pieces at all), a grid that is all one run per ray (only full pieces and one tail)."""
    import nerfacc_b200 as nf

    rng = np.random.default_rng(5 + n_rays)
    res = 32
    o = rng.standard_normal((n_rays, 3)); o = (2.0 * o / np.maximum(np.linalg.norm(o, axis=-1, keepdims=True), 1e-9)).astype(np.float32)
    d = -o + 0.3 * rng.standard_normal((n_rays, 3)); d = (d / np.maximum(np.linalg.norm(d, axis=-1, keepdims=True), 1e-9)).astype(np.float32)
    lin = (np.arange(res) + 0.5) / res * 2 - 1
    X, Y, Z = np.meshgrid(lin, lin, lin, indexing="ij")
    rad = np.sqrt(X * X + Y * Y + Z * Z)
    bins = {"empty": np.zeros((1, res, res, res), bool), "full": np.ones((1, res, res, res), bool),
            "shell": ((rad > 0.5) & (rad < 0.8))[None]}[fill]
    est = nf.OccGridEstimator([-1.0, -1.0, -1.0, 1.0, 1.0, 1.0], resolution=res, levels=1).to(device)
    est.binaries = torch.from_numpy(bins).to(device)
    to, td = torch.from_numpy(o).to(device), torch.from_numpy(d).to(device)
    kw = dict(render_step_size=2e-3, cone_angle=0.003, near_plane=0.1)
    for _ in range(2):  # exact buffers, then the speculative expansion
        ri, ts, te = est.sampling(to, td, **kw)
        if n_rays == 0:
            assert ri.numel() == 0 and ts.numel() == 0
            continue
        r_ri, r_ts, r_te, _ = oracle.sampling(o, d, bins, est.aabbs.cpu().numpy(), **kw)
        assert ri.shape[0] == r_ri.shape[0]
        assert (ri.cpu().numpy() == r_ri).all() and (ts.cpu().numpy() == r_ts).all() and (te.cpu().numpy() == r_te).all()
        if fill == "empty":
            assert ri.numel() == 0
