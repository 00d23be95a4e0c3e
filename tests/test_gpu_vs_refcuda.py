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
from tests._util import packed_info_from_counts, ragged_counts, record

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


def _err(x, t):
    """max |x - t| / (|t| + 1e-2 rms(t)) — the metric of bench_configs.err_vs."""
    from bench_configs import err_vs

    return err_vs(x, t)


def _check(what, ours, ref_out, truth, measured=None):
    """north_star's float bar with numbers behind it: ours within 1e-5 of a float64 evaluation, or
    as close to float64 as the reference's own kernels are (bench_configs.as_close_as_reference)."""
    from bench_configs import as_close_as_reference, rms_err

    e_o, e_r = _err(ours, truth), _err(ref_out, truth)
    r_o, r_r = rms_err(ours, truth), rms_err(ref_out, truth)
    if measured is not None:
        measured[what] = {"max": [e_o, e_r], "rms": [r_o, r_r]}
    assert as_close_as_reference(e_o, e_r, r_o, r_r), (what, "max", e_o, e_r, "rms", r_o, r_r)


def _f64_scan(x64, pi, ri, op, excl):
    """Packed scan in float64 (differentiable): global cumsum minus the ray's offset; products
    through logarithms (inputs are positive)."""
    v = x64 if op == "sum" else torch.log(x64)
    cs = torch.cumsum(v, 0)
    before = (cs - v)[pi[:, 0].clamp_max(max(v.numel() - 1, 0))][ri]  # global prefix at the ray's first item
    out = (cs - v if excl else cs) - before
    return out if op == "sum" else torch.exp(out)


def test_packed_scans_vs_reference_cuda(ref, device):
    """The four packed scans and their gradients at 2^16 ragged rays (incl. a 50 000-sample ray):
    ours and the reference kernels are both measured against a float64 evaluation with the same
    metric; ours must be within 1e-5 (north_star) or as close to float64 as the reference is.
    The measured figures go to DESIGN.md section 2 (via tests/_util.record)."""
    import nerfacc_b200 as nf

    counts = ragged_counts(1 << 16, mean=12, seed=11, long_rays=((5, 50000), (4000, 1)))
    pi = torch.from_numpy(packed_info_from_counts(counts)).to(device)
    ri = torch.repeat_interleave(torch.arange(len(counts), device=device), pi[:, 1])
    n = int(counts.sum())
    g0 = torch.Generator(device=device).manual_seed(3)
    x = torch.rand(n, device=device, generator=g0) * 0.2 + 0.85
    g = torch.rand(n, device=device, generator=g0)
    cases = (("sum", False, nf.inclusive_sum), ("sum", True, nf.exclusive_sum),
             ("prod", False, nf.inclusive_prod), ("prod", True, nf.exclusive_prod))
    measured = {}
    for op, excl, fn in cases:
        xr = x.clone().requires_grad_(True)
        y = fn(xr, pi)
        y_ref = ref.scan(x, pi, op, excl)
        (gx,) = torch.autograd.grad((y * g).sum(), xr)
        if op == "sum":
            gx_ref = ref.scan(g, pi, "sum", excl, backward=True)
        else:
            gx_ref = ref.prod_backward(g, x, y_ref, pi, excl)
        x64 = x.double().requires_grad_(True)
        y64 = _f64_scan(x64, pi, ri, op, excl)
        (gx64,) = torch.autograd.grad((y64 * g.double()).sum(), x64)
        name = f"{'exclusive' if excl else 'inclusive'}_{op}"
        _check(name, y.detach(), y_ref, y64.detach(), measured)
        _check(name + " grad", gx, gx_ref, gx64, measured)
    record("packed scans, 2^16 ragged rays incl. a 50000-sample ray: [ours_vs_f64, ref_cuda_vs_f64]", measured)


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


# ------------------------------------------------------------------ the benchmarked shapes
def test_sampling_cfg2_full_size_vs_reference_cuda(ref, device):
    """BASELINE.json configs[1] at its full 2^20 rays (36.4 M samples): the run pool, the look-back
    chains, the 32-bit cursors and the 32-rays-per-warp batching all behave differently from the
    2^16-ray case above.  Everything the march emits is bit-exact."""
    import nerfacc_b200 as nf
    from bench import AABB, STEP_SIZE, make_grid, make_rays
    from bench_configs import ref_sampling

    n = 1 << 20
    o_np, d_np = make_rays(n, 1234)
    o, d = torch.from_numpy(o_np).to(device), torch.from_numpy(d_np).to(device)
    est = nf.OccGridEstimator(AABB.tolist(), resolution=128, levels=1).to(device)
    est.binaries = torch.from_numpy(make_grid()).to(device)
    ri, ts, te = est.sampling(o, d, render_step_size=STEP_SIZE)
    r_ri, r_ts, r_te, r_pi = ref_sampling(ref, o, d, est.binaries, est.aabbs, 0.0, 1e10, STEP_SIZE)
    assert ri.shape[0] == r_ri.shape[0] > 30 * n
    assert torch.equal(ri, r_ri) and torch.equal(ts, r_ts) and torch.equal(te, r_te)
    assert torch.equal(nf.pack_info(ri, n), r_pi)


def test_cfg3_weights_and_visibility_vs_reference_cuda(ref, device):
    """BASELINE.json configs[2]: ~2^24 ragged samples over 2^20 rays (SURVEY 8(d) count
    distribution, plus one 131 072-sample ray and one 2 048-sample ray): render_weight_from_alpha
    forward + backward within 1e-5 of the reference kernels' composition (and both measured against
    float64), visibility + compaction equal to the reference's mask selects."""
    import nerfacc_b200 as nf
    from bench_configs import cfg3_counts, ref_weight_from_alpha
    from nerfacc_b200 import _occ_ops

    n_rays = 1 << 20
    c = cfg3_counts(n_rays, 1 << 24)
    c[12345], c[777] = 131072, 2048
    cnt = torch.from_numpy(c).to(device)
    ri = torch.repeat_interleave(torch.arange(n_rays, device=device), cnt)
    pi = nf.pack_info(ri, n_rays)
    S = ri.numel()
    assert abs(S - (1 << 24)) < 0.02 * (1 << 24)
    g0 = torch.Generator(device=device).manual_seed(1234)
    alphas = torch.rand(S, device=device, generator=g0) ** 4
    g = torch.rand(S, device=device, generator=g0)
    a1 = alphas.clone().requires_grad_(True)
    w, T = nf.render_weight_from_alpha(a1, packed_info=pi)
    (ga,) = torch.autograd.grad(w, a1, g)
    a2 = alphas.clone().requires_grad_(True)
    w_r, T_r = ref_weight_from_alpha(ref, a2, pi)
    (ga_r,) = torch.autograd.grad(w_r, a2, g)
    # float64 truth: T = exp(segmented exclusive sum of log1p(-a)); the gradient by autograd
    a3 = alphas.double().requires_grad_(True)
    lg = torch.log1p(-a3)
    cs = torch.cumsum(lg, 0) - lg
    T64 = torch.exp(cs - cs[pi[:, 0].clamp_max(S - 1)][ri])
    w64 = T64 * a3
    (ga64,) = torch.autograd.grad(w64, a3, g.double())
    measured = {}
    _check("weights", w.detach(), w_r.detach(), w64.detach(), measured)
    _check("trans", T.detach(), T_r.detach(), T64.detach(), measured)
    # The gradient of the product scan divides by clamp_min(1 - a, 1e-10) (csrc/scan.cu:299-300): where
    # 1 - a is tiny that is not the float64 derivative, so those elements are compared with the
    # reference's own kernels only.
    sane = alphas < 0.999
    _check("grad_alphas", ga[sane], ga_r[sane], ga64[sane], measured)
    e_or = _err(ga, ga_r)
    measured["grad_alphas ours_vs_ref_cuda (all elements, max)"] = e_or
    record("cfg3 render_weight_from_alpha S=2^24: [ours_vs_f64, ref_cuda_vs_f64]", measured)
    assert e_or <= max(1e-5, 4.0 * measured["grad_alphas"]["max"][1]), e_or
    # visibility -> compaction (occ_grid.py:199-220)
    ts = torch.rand(S, device=device, generator=g0)
    te = ts + 0.01
    eps, thre = 1e-4, 1e-2
    o_ri, o_ts, o_te, o_pi, _ = _occ_ops.visibility_compact(False, alphas, ts, te, pi, ri, eps, thre)
    vis = (ref.scan(1.0 - alphas, pi, "prod", True) >= eps) & (alphas >= thre)
    # T rounds differently in the two scans: samples within 1e-6 relative of the threshold may flip
    if not torch.equal(o_ri, ri[vis]):
        T_ours = nf.render_transmittance_from_alpha(alphas, packed_info=pi)
        T_ref = ref.scan(1.0 - alphas, pi, "prod", True)
        flip = ((T_ours >= eps) != (T_ref >= eps))
        assert int(flip.sum()) <= 4 and bool(((T_ref[flip] - eps).abs() <= 1e-6 * eps).all())
    else:
        assert torch.equal(o_ts, ts[vis]) and torch.equal(o_te, te[vis])
        assert torch.equal(o_pi, nf.pack_info(ri[vis].clone(), n_rays))


def test_cfg5_resampling_full_size_vs_reference_cuda(ref, device):
    """BASELINE.json configs[4]: 2^20 rays, edges 2 -> 257 -> 97 -> 49 through importance_sampling
    (centres and edges), then searchsorted of the 49 final edges in the 257: all bit-exact."""
    import nerfacc_b200 as nf

    n_rays = 1 << 20
    g0 = torch.Generator(device=device).manual_seed(1234)
    cd = torch.tensor([0.0, 1.0], device=device).repeat(n_rays, 1)
    lv = cd.clone()
    key = None
    for n_out in (256, 96, 48):
        iv, sm = nf.importance_sampling(nf.RayIntervals(vals=lv), cd, n_out)
        e_ref, c_ref = ref.importance_sampling(lv, cd, n_out)
        assert torch.equal(iv.vals, e_ref) and torch.equal(sm.vals, c_ref), n_out
        iv2, _ = nf.importance_sampling(nf.RayIntervals(vals=lv), cd, n_out, _samples=False)
        assert torch.equal(iv2.vals, e_ref)  # the intervals-only call PropNetEstimator makes
        del e_ref, c_ref, sm, iv2
        lv = iv.vals
        if key is None:
            key = lv
        # the next level's CDF: a bumpy density with empty stretches (degenerate bins included)
        wgt = torch.rand((n_rays, n_out), device=device, generator=g0) ** 3
        wgt = wgt * (torch.rand((n_rays, n_out), device=device, generator=g0) > 0.3)
        cd = torch.cat([torch.zeros((n_rays, 1), device=device), torch.cumsum(wgt + 1e-12, -1)], -1)
        cd = cd / cd[:, -1:]
        del wgt
    l, r = nf.searchsorted(nf.RayIntervals(vals=key), nf.RayIntervals(vals=lv))
    l_ref, r_ref = ref.searchsorted(key, lv)
    assert torch.equal(l, l_ref) and torch.equal(r, r_ref)
