"""nfb_visibility_compact — OccGridEstimator.sampling's "skip invisible space" step
(nerfacc/estimators/occ_grid.py:186-220) as one kernel: visibility -> compaction -> the
survivors' packed_info.

Checked bit for bit (it is index work once T is known, and T comes from the same scan
arithmetic as render_visibility_*, which tests/test_gpu_volrend.py pins to the oracle) against
the reference's own sequence spelled out in torch: mask = render_visibility_*(...);
ray_indices[mask], t_starts[mask], t_ends[mask]; pack_info(ray_indices[mask]).
"""
import numpy as np
import pytest
import torch

import oracle
from tests._util import packed_info_from_counts, ragged_counts, ray_indices_from_counts

pytestmark = pytest.mark.gpu


def _layout(counts, device, seed):
    import nerfacc_b200 as nf

    ri = torch.from_numpy(ray_indices_from_counts(counts)).to(device)
    n_rays = len(counts)
    pi = nf.pack_info(ri, n_rays)  # canonical + dense, as sampling produces it
    assert torch.equal(pi.cpu(), torch.from_numpy(packed_info_from_counts(counts)))
    g = torch.Generator(device=device).manual_seed(seed)
    n = ri.numel()
    ts = torch.rand(n, device=device, generator=g)
    te = ts + 0.01 + 0.05 * torch.rand(n, device=device, generator=g)
    return ri, pi, ts, te, g


CASES = {
    "ragged+long": lambda: ragged_counts(3000, seed=1, long_rays=((5, 9000), (6, 4096), (2999, 700))),
    "exact_tiles": lambda: np.array([4096, 0, 0, 4096, 1, 4095, 0], dtype=np.int64),  # n = 3 tiles exactly
    "trailing_empty": lambda: np.array([0, 0, 7, 0, 300, 0, 0, 0], dtype=np.int64),
    "tiny": lambda: np.array([3], dtype=np.int64),
    "one_long_ray": lambda: np.array([0, 20000, 0], dtype=np.int64),
    "many_short": lambda: ragged_counts(40000, mean=3.0, seed=3),
}


@pytest.mark.parametrize("case", sorted(CASES))
@pytest.mark.parametrize("mode", ["density", "alpha"])
@pytest.mark.parametrize("eps,thre", [(1e-4, 0.0), (1e-2, 0.05), (0.0, 0.3)])
def test_visibility_compact_equals_mask_select_and_pack_info(device, case, mode, eps, thre):
    import nerfacc_b200 as nf
    from nerfacc_b200 import _occ_ops

    counts = CASES[case]()
    ri, pi, ts, te, g = _layout(counts, device, seed=len(case))
    n, n_rays = ri.numel(), len(counts)
    if mode == "density":
        vals = 40.0 * torch.rand(n, device=device, generator=g) ** 3
        mask = nf.render_visibility_from_density(ts, te, vals, packed_info=pi, early_stop_eps=eps, alpha_thre=thre)
    else:
        vals = torch.rand(n, device=device, generator=g) ** 4
        mask = nf.render_visibility_from_alpha(vals, packed_info=pi, early_stop_eps=eps, alpha_thre=thre)
    o_ri, o_ts, o_te, o_pi, o_mask = _occ_ops.visibility_compact(
        mode == "density", vals, ts, te, pi, ri, eps, thre, want_mask=True)
    assert torch.equal(o_mask, mask)
    assert 0 < int(mask.sum()) or case == "tiny"
    assert torch.equal(o_ri, ri[mask]) and torch.equal(o_ts, ts[mask]) and torch.equal(o_te, te[mask])
    # the survivors' packed_info == nerfacc.pack_info of the compacted ray_indices (pack.py:38-46)
    assert torch.equal(o_pi, nf.pack_info(ri[mask].clone(), n_rays))
    # and against the CPU oracle: its sequential fp32 T may round differently from the tile scan,
    # so samples whose T (or alpha) sits within 1e-5 relative of a threshold are left out
    if mode == "alpha" and n < 100000:
        a_np, pi_np = vals.cpu().numpy(), pi.cpu().numpy()
        _, T = oracle.render_weight_from_alpha(a_np, pi_np)
        want = oracle.visibility(T, a_np, eps, thre)
        sure = (np.abs(T - eps) > 1e-5 * eps) & (np.abs(a_np - thre) > 1e-5 * max(thre, 1e-30))
        assert np.array_equal(o_mask.cpu().numpy()[sure], want[sure].astype(bool))


def test_sampling_with_sigma_fn_is_one_kernel_and_keeps_the_fast_layout(device):
    """OccGridEstimator.sampling(sigma_fn=...) -> rendering(ray_indices): after the callback one
    compaction kernel (+ the per-ray count fix-up) runs, the survivors' packed_info travels with
    ray_indices, and rendering lands on the warp-tile path ("bits") without a pack_info pass."""
    import nerfacc_b200 as nf
    from nerfacc_b200 import _C, _packing, _render_ops

    res, n_rays = 64, 20000
    rng = np.random.default_rng(5)
    lin = (np.arange(res) + 0.5) / res * 3.0 - 1.5
    X, Y, Z = np.meshgrid(lin, lin, lin, indexing="ij")
    rad = np.sqrt(X**2 + Y**2 + Z**2)
    bins = ((rad >= 0.8) & (rad <= 0.9))[None]
    o = rng.standard_normal((n_rays, 3))
    o = (4.0 * o / np.linalg.norm(o, axis=-1, keepdims=True)).astype(np.float32)
    d = -o / np.linalg.norm(o, axis=-1, keepdims=True) + 0.1 * rng.standard_normal((n_rays, 3)).astype(np.float32)
    d = (d / np.linalg.norm(d, axis=-1, keepdims=True)).astype(np.float32)
    est = nf.OccGridEstimator([-1.5] * 3 + [1.5] * 3, resolution=res).to(device)
    est.binaries = torch.from_numpy(bins).to(device)
    est.occs = est.binaries.flatten().float()
    to, td = torch.from_numpy(o).to(device), torch.from_numpy(d).to(device)

    def sigma_fn(t_starts, t_ends, ray_indices):  # examples/utils.py:82-95 pattern
        pos = to[ray_indices] + td[ray_indices] * ((t_starts + t_ends) * 0.5)[:, None]
        return 50.0 * ((pos.norm(dim=-1) >= 0.8) & (pos.norm(dim=-1) <= 0.9)).float()

    ri0, ts0, te0 = est.sampling(to, td, render_step_size=5e-3)
    sig0 = sigma_fn(ts0, te0, ri0)
    mask = nf.render_visibility_from_density(ts0, te0, sig0, ray_indices=ri0, n_rays=n_rays, early_stop_eps=1e-4)
    _C.prof_enable(True)  # records which C-ABI entry points run
    try:
        ri, ts, te = est.sampling(to, td, sigma_fn=sigma_fn, render_step_size=5e-3, early_stop_eps=1e-4)
        called = _C.prof_collect()
    finally:
        _C.prof_enable(False)
    # march (count, expand), the user's callback (torch), then ONE native call: no mask kernel,
    # no separate compaction, no pack_info
    # (traverse_expand: the second and later calls expand the runs ahead of the count read-back)
    assert sorted(called) in (["traverse_count", "traverse_fill_sampling", "visibility_compact"],
                              ["traverse_count", "traverse_expand", "visibility_compact"]), called
    assert called["visibility_compact"][0] == 1
    assert 0 < ri.numel() < ri0.numel()
    assert torch.equal(ri, ri0[mask]) and torch.equal(ts, ts0[mask]) and torch.equal(te, te0[mask])
    pi = _packing.recall_packed_info(ri, n_rays)
    assert pi is not None and torch.equal(pi, nf.pack_info(ri.clone(), n_rays))
    assert _render_ops._render_mode(ri.numel(), n_rays, pi) == "bits"
