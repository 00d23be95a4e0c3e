"""The strongest parity check: this library against the reference's OWN CUDA kernels
(oracle/_ref/nerfacc_cuda.so, the unmodified extension compiled for sm_100a) on the same
device tensors, at sizes well beyond what the CPU oracle is asked to do.

Bit-exact for everything the traversal emits; 1e-5 relative (+ a magnitude-scaled atol) for the
float scans, weights and gradients, as north_star states.  Skipped when the reference
extension has not been built (oracle/build_ref.py).
"""
import numpy as np
import pytest
import torch

from tests import _refcuda as R
from tests._util import packed_info_from_counts, ragged_counts

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ref(device):
    if not R.available():
        pytest.skip("oracle/_ref/nerfacc_cuda.so not built")
    return R


def _close(a, b, rtol=1e-5, atol_scale=1e-6):
    a, b = a.double(), b.double()
    atol = atol_scale * float(b.abs().max().clamp_min(1e-30))
    bad = (a - b).abs() > atol + rtol * b.abs()
    assert not bad.any(), f"{int(bad.sum())}/{bad.numel()} mismatches, max abs err {float((a - b).abs().max()):.3e}"


def test_packed_scans_vs_reference_cuda(ref, device):
    import nerfacc_b200 as nf

    counts = ragged_counts(1 << 16, mean=12, seed=11, long_rays=((5, 50000), (4000, 1)))
    pi = torch.from_numpy(packed_info_from_counts(counts)).to(device)
    n = int(counts.sum())
    g0 = torch.Generator(device=device).manual_seed(3)
    x = torch.rand(n, device=device, generator=g0) * 0.2 + 0.85
    g = torch.rand(n, device=device, generator=g0)
    cases = (("sum", False, nf.inclusive_sum), ("sum", True, nf.exclusive_sum),
             ("prod", False, nf.inclusive_prod), ("prod", True, nf.exclusive_prod))
    for op, excl, fn in cases:
        xr = x.clone().requires_grad_(True)
        y = fn(xr, pi)
        y_ref = ref.scan(x, pi, op, excl)
        # sums of ~50000 terms: both sides round differently; compare at the row's magnitude
        _close(y, y_ref, rtol=2e-5 if op == "sum" else 1e-4, atol_scale=1e-6)
        (gx,) = torch.autograd.grad((y * g).sum(), xr)
        if op == "sum":
            gx_ref = ref.scan(g, pi, "sum", excl, backward=True)
        else:
            gx_ref = ref.prod_backward(g, x, y_ref, pi, excl)
        _close(gx, gx_ref, rtol=1e-4, atol_scale=2e-6)


@pytest.mark.parametrize("kw", [
    dict(step_size=1e-3), dict(step_size=0.013, cone_angle=0.004), dict(step_size=0.0),
    dict(step_size=2e-2, limit=37),
])
def test_traverse_grids_vs_reference_cuda(ref, device, kw):
    import nerfacc_b200 as nf

    rng = np.random.default_rng(42)
    n, m, res = 4096, 4, 32
    o = torch.from_numpy(rng.standard_normal((n, 3)).astype(np.float32)).to(device)
    d = rng.standard_normal((n, 3)).astype(np.float32)
    d = torch.from_numpy(d / np.linalg.norm(d, axis=-1, keepdims=True)).to(device)
    bins = torch.from_numpy(rng.random((m, res, res, res)) < 0.3).to(device)
    base = torch.tensor([-1.0, -1, -1, 1, 1, 1], device=device)
    aabbs = torch.stack([base * 2**i for i in range(m)])
    near = torch.rand(n, device=device) * 0.2
    far = torch.full((n,), 6.0, device=device)
    iv_r, sm_r, term_r = ref.traverse_grids(o, d, bins, aabbs, near, far, kw.get("step_size"),
                                            kw.get("cone_angle", 0.0), kw.get("limit"))
    iv, sm, term = nf.traverse_grids(o, d, bins, aabbs, near_planes=near, far_planes=far,
                                     step_size=kw["step_size"], cone_angle=kw.get("cone_angle", 0.0),
                                     traverse_steps_limit=kw.get("limit"))
    assert sm.vals.shape[0] == sm_r["vals"].shape[0] > 10000
    for k, v in iv_r.items():
        assert torch.equal(getattr(iv, k), v), ("intervals", k)
    for k, v in sm_r.items():
        assert torch.equal(getattr(sm, k), v), ("samples", k)
    has = sm_r["packed_info"][:, 1] > 0
    assert torch.equal(term[has], term_r[has])


def test_sampling_cfg2_vs_reference_cuda(ref, device):
    """BASELINE.json configs[1] at 2^16 rays: ray_indices, t_starts, t_ends bit-exact."""
    import nerfacc_b200 as nf
    from bench import AABB, STEP_SIZE, make_grid, make_rays

    n = 1 << 16
    o_np, d_np = make_rays(n, 99)
    o, d = torch.from_numpy(o_np).to(device), torch.from_numpy(d_np).to(device)
    bins = torch.from_numpy(make_grid()).to(device)
    est = nf.OccGridEstimator(AABB.tolist(), resolution=128, levels=1).to(device)
    est.binaries = bins
    ri, ts, te = est.sampling(o, d, render_step_size=STEP_SIZE)
    near = torch.zeros(n, device=device)
    far = torch.full((n,), 1e10, device=device)
    iv_r, sm_r, _ = ref.traverse_grids(o, d, bins, est.aabbs, near, far, STEP_SIZE, 0.0)
    assert torch.equal(ri, sm_r["ray_indices"])
    assert torch.equal(ts, iv_r["vals"][iv_r["is_left"]])
    assert torch.equal(te, iv_r["vals"][iv_r["is_right"]])
    assert torch.equal(nf.pack_info(ri, n), sm_r["packed_info"])
    assert ri.shape[0] > 30 * n * 0.9


def test_ray_aabb_vs_reference_cuda(ref, device):
    import nerfacc_b200 as nf

    g0 = torch.Generator(device=device).manual_seed(5)
    o = torch.rand((1 << 16, 3), device=device, generator=g0)
    d = torch.randn((1 << 16, 3), device=device, generator=g0)
    d = d / d.norm(dim=-1, keepdim=True)
    lo = torch.rand((13, 3), device=device, generator=g0)
    bb = torch.cat([lo, lo + torch.rand((13, 3), device=device, generator=g0)], -1)
    for args in ((), (0.1, 0.7, -1.0)):
        a = nf.ray_aabb_intersect(o, d, bb, *args)
        b = ref.ray_aabb_intersect(o, d, bb, *args)
        for x, y in zip(a, b):
            assert torch.equal(x, y)


def test_pdf_vs_reference_cuda(ref, device):
    """importance_sampling (batched + packed input) and searchsorted at 2^15 rays."""
    import nerfacc_b200 as nf

    g0 = torch.Generator(device=device).manual_seed(17)
    n_rays, n_in, n_out = 1 << 15, 65, 48
    vals = torch.sort(torch.rand((n_rays, n_in), device=device, generator=g0) * 8 + 0.1, dim=-1).values
    w = torch.rand((n_rays, n_in - 1), device=device, generator=g0)
    w = w * (torch.rand((n_rays, n_in - 1), device=device, generator=g0) > 0.2) + 1e-4
    cdf = torch.cat([torch.zeros((n_rays, 1), device=device), torch.cumsum(w, -1)], -1)
    cdf = cdf / cdf[:, -1:]
    iv, sm = nf.importance_sampling(nf.RayIntervals(vals=vals), cdf, n_out)
    e_ref, c_ref = ref.importance_sampling(vals, cdf, n_out)
    assert torch.equal(sm.vals, c_ref)
    assert torch.equal(iv.vals, e_ref)
    q = torch.rand((n_rays, 37), device=device, generator=g0) * 9
    l, r = nf.searchsorted(nf.RayIntervals(vals=vals), nf.RayIntervals(vals=q))
    l_ref, r_ref = ref.searchsorted(vals, q)
    assert torch.equal(l, l_ref) and torch.equal(r, r_ref)
    # packed key + packed query
    kc = torch.randint(1, 80, (n_rays,), device=device, generator=g0)
    qc = torch.randint(0, 40, (n_rays,), device=device, generator=g0)
    ki = torch.stack([torch.cumsum(kc, 0) - kc, kc], -1)
    qi = torch.stack([torch.cumsum(qc, 0) - qc, qc], -1)
    kv = torch.rand(int(kc.sum()), device=device, generator=g0)
    rid = torch.repeat_interleave(torch.arange(n_rays, device=device), kc)
    kv = torch.sort(kv + rid.float() * 2)[0] - rid.float() * 2  # sorted within each ray
    qv = torch.rand(int(qc.sum()), device=device, generator=g0) * 1.2 - 0.1
    l, r = nf.searchsorted(nf.RaySamples(vals=kv, packed_info=ki), nf.RaySamples(vals=qv, packed_info=qi))
    l_ref, r_ref = ref.searchsorted(kv, qv, ki, qi)
    assert torch.equal(l, l_ref) and torch.equal(r, r_ref)
