"""importance_sampling / searchsorted / PropNetEstimator on the GPU vs the CPU oracle, the golden
fixtures (reference CUDA kernels on B200: golden_gpu `pdf/*`; reference torch CPU path:
golden_cpu `pdf/*`, `loss/*`, `stot/*`) and the reference's own documented examples
(nerfacc/pdf.py:40-56, 108-120; tests/test_pdf.py).

Positions: bit-exact vs the oracle (same rounding sequence), 1e-5 rel vs the stored GPU outputs;
indices: exact.
"""
import os

import numpy as np
import pytest
import torch

import oracle
from tests._util import packed_info_from_counts

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _cdf_rows(rng, n_rays, n_edges, flat_share=0.2):
    """Sorted edges in [0.1, 9] and a CDF with some exactly-flat (zero-mass) bins."""
    vals = np.sort(rng.uniform(0.1, 9.0, (n_rays, n_edges)).astype(np.float32), axis=-1)
    w = rng.random((n_rays, n_edges - 1)).astype(np.float32)
    w[rng.random(w.shape) < flat_share] = 0.0
    w[:, 0] += 1e-3
    cdf = np.concatenate([np.zeros((n_rays, 1), np.float32), np.cumsum(w, -1, dtype=np.float32)], -1)
    cdf = (cdf / cdf[:, -1:]).astype(np.float32)
    return vals, cdf


def _t(a, device):
    return torch.from_numpy(np.ascontiguousarray(a)).to(device)


@pytest.mark.parametrize("n_in,n_out", [(2, 1), (2, 64), (65, 48), (129, 1), (33, 257), (3000, 500)])
@pytest.mark.parametrize("stratified", [False, True])
def test_importance_sampling_batched_vs_oracle(device, n_in, n_out, stratified):
    import nerfacc_b200 as nf

    rng = np.random.default_rng(n_in * 1000 + n_out)
    n_rays = 257
    vals, cdf = _cdf_rows(rng, n_rays, n_in)
    bias = (1.0 - rng.random(n_rays)).astype(np.float32) if stratified else None
    iv, sm = nf.importance_sampling(nf.RayIntervals(vals=_t(vals, device)), _t(cdf, device), n_out,
                                    stratified, _bias=None if bias is None else _t(bias, device))
    e_ref, c_ref = oracle.importance_sampling(vals, cdf, n_out, bias=bias)
    assert iv.vals.shape == (n_rays, n_out + 1) and sm.vals.shape == (n_rays, n_out)
    assert np.array_equal(sm.vals.cpu().numpy(), c_ref)
    assert np.array_equal(iv.vals.cpu().numpy(), e_ref)
    assert iv.packed_info is None and sm.packed_info is None


@pytest.mark.parametrize("n_rays", [1, 3, 70])
@pytest.mark.parametrize("n_in,n_out", [(97, 48), (5, 80), (2, 256), (257, 96)])
def test_importance_sampling_batched_tile_edges(device, n_rays, n_in, n_out):
    """The tile kernel's variants at ray counts that leave tiles, warps and half-warps partly empty:
    a ray per half-warp (48 samples), the per-ray slope table (5 -> 80), the single bin (2 -> 256), the
    ladder search (257 -> 96); spans too short or misaligned for the bulk copy take the plain loop."""
    import nerfacc_b200 as nf

    rng = np.random.default_rng(n_rays * 131 + n_in)
    vals, cdf = _cdf_rows(rng, n_rays, n_in)
    for with_centres in (True, False):
        iv, sm = nf.importance_sampling(nf.RayIntervals(vals=_t(vals, device)), _t(cdf, device), n_out,
                                        _samples=with_centres)
        e_ref, c_ref = oracle.importance_sampling(vals, cdf, n_out)
        assert np.array_equal(iv.vals.cpu().numpy(), e_ref)
        if with_centres:
            assert np.array_equal(sm.vals.cpu().numpy(), c_ref)


def test_importance_sampling_packed_input_vs_oracle(device):
    import nerfacc_b200 as nf

    rng = np.random.default_rng(7)
    counts = rng.integers(2, 90, 300)
    counts[5] = 5000  # longer than the shared-memory staging: global-memory search
    pi = packed_info_from_counts(counts)
    vals = np.concatenate([np.sort(rng.uniform(0, 5, c)) for c in counts]).astype(np.float32)
    cdf = np.concatenate([np.linspace(0, 1, c) ** 2 for c in counts]).astype(np.float32)
    iv, sm = nf.importance_sampling(
        nf.RayIntervals(vals=_t(vals, device), packed_info=_t(pi, device)), _t(cdf, device), 40)
    e_ref, c_ref = oracle.importance_sampling(vals, cdf, 40, packed_info=pi)
    assert np.array_equal(sm.vals.cpu().numpy(), c_ref)
    assert np.array_equal(iv.vals.cpu().numpy(), e_ref)


def test_importance_sampling_per_ray_counts(device):
    """Tensor n_intervals_per_ray: flattened outputs; each ray equals the int-n result on it."""
    import nerfacc_b200 as nf

    rng = np.random.default_rng(8)
    n_rays, n_in = 64, 33
    vals, cdf = _cdf_rows(rng, n_rays, n_in)
    n_per = rng.integers(0, 70, n_rays)
    n_per[:4] = (0, 1, 2, 0)
    iv, sm = nf.importance_sampling(nf.RayIntervals(vals=_t(vals, device)), _t(cdf, device),
                                    _t(n_per, device))
    e_cnt = (n_per + 1) * (n_per > 0)
    assert np.array_equal(sm.packed_info.cpu().numpy(), packed_info_from_counts(n_per))
    assert np.array_equal(iv.packed_info.cpu().numpy(), packed_info_from_counts(e_cnt))
    assert np.array_equal(sm.ray_indices.cpu().numpy(), np.repeat(np.arange(n_rays), n_per))
    assert np.array_equal(iv.ray_indices.cpu().numpy(), np.repeat(np.arange(n_rays), e_cnt))
    ev, cv = iv.vals.cpu().numpy(), sm.vals.cpu().numpy()
    il, ir = iv.is_left.cpu().numpy(), iv.is_right.cpu().numpy()
    e0 = s0 = 0
    for r in range(n_rays):
        n = int(n_per[r])
        if n == 0:
            continue
        e_ref, c_ref = oracle.importance_sampling(vals[r:r + 1], cdf[r:r + 1], n)
        assert np.array_equal(cv[s0:s0 + n], c_ref[0]), r
        assert np.array_equal(ev[e0:e0 + n + 1], e_ref[0]), r
        assert il[e0:e0 + n + 1].tolist() == [True] * n + [False]
        assert ir[e0:e0 + n + 1].tolist() == [False] + [True] * n
        e0 += n + 1
        s0 += n
    assert e0 == ev.shape[0] and s0 == cv.shape[0]


def test_importance_sampling_doc_example_and_empty(device):
    import nerfacc_b200 as nf

    iv = nf.RayIntervals(vals=torch.tensor([0.0, 1.0, 0.0, 1.0, 2.0], device=device),
                         packed_info=torch.tensor([[0, 2], [2, 3]], device=device))
    cdfs = torch.tensor([0.0, 0.5, 0.0, 0.5, 1.0], device=device)
    out_iv, out_sm = nf.importance_sampling(iv, cdfs, 2)
    assert out_iv.vals.tolist() == [[0.0, 0.5, 1.0], [0.0, 1.0, 2.0]]
    assert out_sm.vals.tolist() == [[0.25, 0.75], [0.5, 1.5]]
    e = nf.RayIntervals(vals=torch.zeros((0, 5), device=device))
    out_iv, out_sm = nf.importance_sampling(e, torch.zeros((0, 5), device=device), 7)
    assert out_iv.vals.shape == (0, 8) and out_sm.vals.shape == (0, 7)
    with pytest.raises(RuntimeError):
        nf.importance_sampling(nf.RayIntervals(vals=torch.zeros(2, 3)), torch.zeros(2, 3), 4)


def test_stratified_draws_one_offset_per_ray(device):
    import nerfacc_b200 as nf

    n_rays, n_in, n_out = 512, 2, 16
    vals = torch.tensor([0.0, 1.0], device=device).repeat(n_rays, 1)
    torch.manual_seed(3)
    _, a = nf.importance_sampling(nf.RayIntervals(vals=vals), vals.clone(), n_out, True)
    torch.manual_seed(3)
    _, b = nf.importance_sampling(nf.RayIntervals(vals=vals), vals.clone(), n_out, True)
    assert torch.equal(a.vals, b.vals)
    # uniform CDF: centre k = (k + bias)/n with the ray's own bias in (0, 1]
    bias = a.vals[:, 0] * n_out
    assert float(bias.min()) > 0 and float(bias.max()) <= 1 + 1e-6 and float(bias.std()) > 0.2
    k = torch.arange(n_out, device=device)
    torch.testing.assert_close(a.vals, (k[None] + bias[:, None]) / n_out, rtol=1e-5, atol=1e-6)


def test_searchsorted_batched_and_packed_vs_oracle(device):
    import nerfacc_b200 as nf

    rng = np.random.default_rng(9)
    key = np.sort(rng.uniform(0, 1, (130, 77)).astype(np.float32), -1)
    q = rng.uniform(-0.1, 1.1, (130, 45)).astype(np.float32)
    q[:, :5] = key[:, 10:15]  # exact hits: "right" side
    l, r = nf.searchsorted(nf.RayIntervals(vals=_t(key, device)), nf.RayIntervals(vals=_t(q, device)))
    l_ref, r_ref = oracle.searchsorted(key, q)
    assert l.dtype == torch.long and l.shape == q.shape
    assert np.array_equal(l.cpu().numpy(), l_ref) and np.array_equal(r.cpu().numpy(), r_ref)

    kc = rng.integers(1, 60, 200)
    kc[3] = 7000
    qc = rng.integers(0, 30, 200)
    ki, qi = packed_info_from_counts(kc), packed_info_from_counts(qc)
    kf = np.concatenate([np.sort(rng.uniform(0, 1, c)) for c in kc]).astype(np.float32)
    qf = rng.uniform(-0.1, 1.1, int(qc.sum())).astype(np.float32)
    l, r = nf.searchsorted(nf.RaySamples(vals=_t(kf, device), packed_info=_t(ki, device)),
                           nf.RaySamples(vals=_t(qf, device), packed_info=_t(qi, device)))
    l_ref, r_ref = oracle.searchsorted(kf, qf, ki, qi)
    assert np.array_equal(l.cpu().numpy(), l_ref) and np.array_equal(r.cpu().numpy(), r_ref)


def test_searchsorted_doc_example(device):
    import nerfacc_b200 as nf

    seq = nf.RayIntervals(vals=torch.tensor([0.0, 1.0, 0.0, 1.0, 2.0], device=device),
                          packed_info=torch.tensor([[0, 2], [2, 3]], device=device))
    q = nf.RayIntervals(vals=torch.tensor([0.5, 1.5, 2.5], device=device),
                        packed_info=torch.tensor([[0, 1], [1, 2]], device=device))
    l, r = nf.searchsorted(seq, q)
    assert l.tolist() == [0, 3, 3] and r.tolist() == [1, 4, 4]


def test_golden_reference_cuda_pdf(device):
    """Inputs/outputs recorded from the reference's own CUDA kernels on B200."""
    import nerfacc_b200 as nf

    G = np.load(os.path.join(GOLD, "golden_gpu.npz"))
    iv, sm = nf.importance_sampling(nf.RayIntervals(vals=_t(G["pdf/vals"], device)),
                                    _t(G["pdf/cdfs"], device), G["pdf/centers"].shape[1])
    np.testing.assert_allclose(sm.vals.cpu().numpy(), G["pdf/centers"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(iv.vals.cpu().numpy(), G["pdf/edges"], rtol=1e-5, atol=1e-6)
    l, r = nf.searchsorted(nf.RayIntervals(vals=_t(G["pdf/vals"], device)),
                           nf.RayIntervals(vals=_t(G["pdf/query"], device)))
    assert np.array_equal(l.cpu().numpy(), G["pdf/ids_left"])
    assert np.array_equal(r.cpu().numpy(), G["pdf/ids_right"])


def test_golden_reference_torch_pdf_loss_stot(device):
    """Reference CPU-side results (its _sample_from_weighted, _lossfun_outer, _transform_stot);
    1e-4 absolute as in the reference's own tests (tests/test_pdf.py:93-94, 126-127)."""
    import nerfacc_b200 as nf
    from nerfacc_b200.estimators.prop_net import _pdf_loss, _transform_stot

    G = np.load(os.path.join(GOLD, "golden_cpu.npz"))
    iv, sm = nf.importance_sampling(nf.RayIntervals(vals=_t(G["pdf/bins"], device)),
                                    _t(G["pdf/cdfs"], device), G["pdf/centers"].shape[1])
    np.testing.assert_allclose(sm.vals.cpu().numpy(), G["pdf/centers"], atol=1e-4)
    np.testing.assert_allclose(iv.vals.cpu().numpy(), G["pdf/edges"], atol=1e-4)

    def cdf_of(w):
        c = torch.cumsum(w, -1)
        return torch.cat([torch.zeros_like(c[:, :1]), c], -1)

    t_q, w_q = _t(G["loss/t_q"], device), _t(G["loss/w_q"], device)
    t_k, w_k = _t(G["loss/t_k"], device), _t(G["loss/w_k"], device)
    loss = _pdf_loss(nf.RayIntervals(vals=t_q), cdf_of(w_q), nf.RayIntervals(vals=t_k), cdf_of(w_k))
    np.testing.assert_allclose(loss.cpu().numpy(), G["loss/outer"], atol=1e-4)

    s = _t(G["stot/s"], device)
    for kind in ("lindisp", "uniform"):
        t = _transform_stot(kind, s, 0.2, 1e3)
        np.testing.assert_allclose(t.cpu().numpy(), G["stot/" + kind], rtol=2e-6, atol=1e-6)


# ----------------------------------------------------------------------------- PropNetEstimator
class _Blob(torch.nn.Module):
    """A tiny density 'network' along the ray: a learnable bump in t."""

    def __init__(self, centre, width):
        super().__init__()
        self.centre = torch.nn.Parameter(torch.tensor(float(centre)))
        self.width = torch.nn.Parameter(torch.tensor(float(width)))

    def forward(self, t_starts, t_ends):
        t = (t_starts + t_ends) * 0.5
        return 5.0 * torch.exp(-((t - self.centre) / self.width) ** 2)


def _oracle_prop_sampling(levels, prop_samples, num_samples, n_rays, near, far, kind):
    """PropNetEstimator.sampling (prop_net.py:76-136) restated on the oracle, unstratified."""
    from nerfacc_b200.estimators.prop_net import _transform_stot

    cdfs = np.tile(np.array([[0.0, 1.0]], np.float32), (n_rays, 1))
    vals = cdfs.copy()
    for fn, n in zip(levels, prop_samples):
        vals, _ = oracle.importance_sampling(vals, cdfs, n)
        t = _transform_stot(kind, torch.from_numpy(vals), near, far)
        ts, te = t[:, :-1], t[:, 1:]
        with torch.no_grad():
            sig = fn(ts, te)
        trans = torch.exp(-torch.cumsum(torch.cat([torch.zeros_like(sig[:, :1]), (sig * (te - ts))[:, :-1]], -1), -1))
        cdfs = (1.0 - torch.cat([trans, torch.zeros_like(trans[:, :1])], -1)).numpy()
    vals, _ = oracle.importance_sampling(vals, cdfs, num_samples)
    t = _transform_stot(kind, torch.from_numpy(vals), near, far)
    return t[:, :-1].numpy(), t[:, 1:].numpy()


@pytest.mark.parametrize("kind", ["lindisp", "uniform"])
def test_propnet_sampling_vs_oracle(device, kind):
    import nerfacc_b200 as nf

    n_rays = 300
    nets = [_Blob(2.0, 0.8), _Blob(2.2, 0.4)]
    est = nf.PropNetEstimator().to(device)
    gpu_nets = [_Blob(2.0, 0.8).to(device), _Blob(2.2, 0.4).to(device)]
    ts, te = est.sampling(gpu_nets, [48, 24], 16, n_rays, 0.2, 6.0, sampling_type=kind)
    assert ts.shape == te.shape == (n_rays, 16) and not ts.requires_grad
    assert len(est.prop_cache) == 0
    ts_ref, te_ref = _oracle_prop_sampling(nets, [48, 24], 16, n_rays, 0.2, 6.0, kind)
    # the network's exp() differs by ulps between CPU and GPU; the resampling amplifies that a bit
    from tests._util import record
    record(f"PropNetEstimator.sampling vs the oracle pipeline on CPU networks ({kind}), 300 rays",
           {"max_rel_ts": float(np.max(np.abs(ts.cpu().numpy() - ts_ref) / np.abs(ts_ref))),
            "max_rel_te": float(np.max(np.abs(te.cpu().numpy() - te_ref) / np.abs(te_ref)))})
    np.testing.assert_allclose(ts.cpu().numpy(), ts_ref, rtol=2e-4, atol=2e-5)
    np.testing.assert_allclose(te.cpu().numpy(), te_ref, rtol=2e-4, atol=2e-5)
    assert bool((te >= ts).all()) and float(ts.min()) >= 0.2 - 1e-5 and float(te.max()) <= 6.0 + 1e-4
    # mass concentrates around the bump
    assert float(((ts > 1.0) & (ts < 3.4)).float().mean()) > 0.8


def test_propnet_loss_and_gradients_vs_torch(device):
    """compute_loss through searchsorted + the fused transmittance backward equals a plain-torch
    restatement (torch.searchsorted envelope loss, exp(-cumsum) transmittance): value and the
    gradients reaching the proposal network's parameters."""
    import nerfacc_b200 as nf
    from nerfacc_b200.estimators.prop_net import _lossfun_outer, _transform_stot

    n_rays = 256
    props = [_Blob(3.0, 1.5).to(device), _Blob(2.5, 1.0).to(device)]
    truth = _Blob(2.0, 0.3).to(device)
    params = [p for m in props for p in m.parameters()]
    est = nf.PropNetEstimator().to(device)
    ts, te = est.sampling(props, [48, 32], 24, n_rays, 0.2, 6.0, sampling_type="uniform",
                          requires_grad=True)
    assert len(est.prop_cache) == 3
    cache = list(est.prop_cache)
    with torch.no_grad():
        trans, _ = nf.render_transmittance_from_density(ts, te, truth(ts, te))
    loss = est.compute_loss(trans, loss_scaler=2.0)
    assert len(est.prop_cache) == 0
    grads = torch.autograd.grad(loss, params)

    def torch_cdf(t, net):
        a, b = t[:, :-1], t[:, 1:]
        sd = net(a, b) * (b - a)
        tr = torch.exp(-(torch.cumsum(sd, -1) - sd))
        return 1.0 - torch.cat([tr, torch.zeros_like(tr[:, :1])], -1)

    s_final = cache[-1][0].vals
    c_final = (1.0 - torch.cat([trans, torch.zeros_like(trans[:, :1])], -1))
    w_final = c_final[:, 1:] - c_final[:, :-1]
    ref = 0.0
    for (iv, _), net in zip(cache[:-1], props):
        c = torch_cdf(_transform_stot("uniform", iv.vals, 0.2, 6.0), net)
        # eps differs (1e-7 in _pdf_loss, finfo.eps in _lossfun_outer): negligible at w ~ 1e-2
        ref = ref + _lossfun_outer(s_final, w_final, iv.vals, c[:, 1:] - c[:, :-1]).mean()
    ref = ref * 2.0
    grads_ref = torch.autograd.grad(ref, params)
    from tests._util import record
    record("PropNetEstimator loss/grads vs the torch.searchsorted restatement (_lossfun_outer), 256 rays",
           {"rel_loss": float((loss - ref).abs() / ref.abs()),
            "rel_grads": [float((g - gr).abs() / gr.abs().clamp_min(1e-30)) for g, gr in zip(grads, grads_ref)],
            "grads_ref": [float(gr) for gr in grads_ref]})
    torch.testing.assert_close(loss, ref, rtol=1e-4, atol=1e-8)
    for g, gr in zip(grads, grads_ref):
        torch.testing.assert_close(g, gr, rtol=2e-3, atol=1e-7)
    assert any(float(g.abs()) > 0 for g in grads)


def test_propnet_update_protocol(device):
    """sampling -> render -> update_every_n_steps, as the reference's training scripts drive it."""
    import nerfacc_b200 as nf
    from nerfacc_b200.estimators.prop_net import get_proposal_requires_grad_fn

    torch.manual_seed(0)
    n_rays = 256
    prop = _Blob(3.0, 1.5).to(device)
    truth = _Blob(2.0, 0.3).to(device)
    opt = torch.optim.SGD(prop.parameters(), lr=1.0)
    sched = torch.optim.lr_scheduler.StepLR(opt, step_size=1, gamma=0.5)
    est = nf.PropNetEstimator(opt, sched).to(device)
    before = [float(p.detach()) for p in prop.parameters()]
    for i in range(4):
        ts, te = est.sampling([prop], [32], 32, n_rays, 0.2, 6.0, sampling_type="lindisp",
                              stratified=True, requires_grad=True)
        assert len(est.prop_cache) == 2
        with torch.no_grad():
            trans, _ = nf.render_transmittance_from_density(ts, te, truth(ts, te))
        loss = est.update_every_n_steps(trans, requires_grad=True)
        assert isinstance(loss, float) and np.isfinite(loss) and loss > 0
        assert len(est.prop_cache) == 0
    assert [float(p.detach()) for p in prop.parameters()] != before
    assert abs(opt.param_groups[0]["lr"] - 1.0 / 16) < 1e-9
    assert est.update_every_n_steps(trans, requires_grad=False) == 0.0
    assert abs(opt.param_groups[0]["lr"] - 1.0 / 32) < 1e-9
    assert float(est.compute_loss(trans)) == 0.0  # empty cache
    with pytest.raises(AssertionError):
        est.update_every_n_steps(trans, requires_grad=True)  # nothing cached
    fn = get_proposal_requires_grad_fn(target=5.0, num_steps=10)
    fired = [fn(s) for s in range(40)]
    assert fired[:6] == [False, True, False, True, False, False] and sum(fired[20:]) == 3


def test_propnet_vs_reference_package_cfg5(device):
    """PropNetEstimator against the reference's OWN PropNetEstimator (baseline/_ref on
    oracle/_ref's kernels) at BASELINE cfg5's shape: 2^20 rays, proposal levels 256 then 96,
    48 final samples, lindisp, analytic proposal densities (two bumps per ray, one trainable
    scale per level); then the proposal loss and its gradients at 2^18 rays.

    The level-1 edges come out of importance_sampling on the uniform CDF: bit-exact.  After that
    the two sides differ in how the batched transmittance is evaluated (the reference:
    exp(-cumsum), here one fused kernel), a few ulps in the CDF, which the inverse-CDF lookup
    amplifies where the density is flat — so positions are compared in the normalised coordinate
    s in [0, 1] (before lindisp stretches them) with an absolute bound, and the figures measured
    here are recorded for DESIGN.md."""
    from tests import _refpkg
    from tests._util import record

    if not _refpkg.available():
        pytest.skip("baseline/_ref (reference package) or oracle/_ref not present")
    import nerfacc_b200 as nf

    ref = _refpkg.load()
    near, far = 0.2, 1e3

    def make_fns(n_rays):
        g = torch.Generator(device=device).manual_seed(1234)
        c1 = 1.0 + 3.0 * torch.rand((n_rays, 1), device=device, generator=g)
        c2 = 5.0 + 20.0 * torch.rand((n_rays, 1), device=device, generator=g)
        scales = [torch.nn.Parameter(torch.tensor(v, device=device)) for v in (1.0, 0.7)]

        def mk(scale):
            def fn(t_starts, t_ends):
                t = (t_starts + t_ends) * 0.5
                return scale * (8.0 * torch.exp(-((t - c1) / 0.3) ** 2) + 0.5 * torch.exp(-((t - c2) / 3.0) ** 2))
            return fn
        return [mk(s) for s in scales], scales, (c1, c2)

    def to_s(t):  # inverse of _transform_stot("lindisp")
        return (1.0 / t - 1.0 / near) / (1.0 / far - 1.0 / near)

    # ---- sampling at the full cfg5 size
    n_rays = 1 << 20
    fns, _, _ = make_fns(n_rays)
    ours, theirs = nf.PropNetEstimator().to(device), ref.PropNetEstimator().to(device)
    ts, te = ours.sampling(fns, [256, 96], 48, n_rays, near, far, "lindisp", False, False)
    ts_r, te_r = theirs.sampling(fns, [256, 96], 48, n_rays, near, far, "lindisp", False, False)
    assert ts.shape == ts_r.shape == (n_rays, 48)
    ds = torch.maximum((to_s(ts) - to_s(ts_r)).abs(), (to_s(te) - to_s(te_r)).abs())
    rel_t = torch.maximum((ts - ts_r).abs() / ts_r, (te - te_r).abs() / te_r)
    m = {"rays": n_rays, "max_abs_ds": float(ds.max()), "frac_ds_le_1e-6": float((ds <= 1e-6).float().mean()),
         "max_rel_dt": float(rel_t.max()), "frac_rel_dt_le_1e-5": float((rel_t <= 1e-5).float().mean()),
         "frac_bit_equal": float(((ts == ts_r) & (te == te_r)).float().mean())}
    record("PropNetEstimator.sampling cfg5 vs the reference's PropNetEstimator", m)
    # measured on a B200 (profiles/r02_*_measured_errors.json): 98.1 % of the edges bit-equal, 99.98 %
    # within 1e-5 relative, the rest sit in flat stretches of the CDF; max |ds| = 3.6e-6
    assert m["frac_rel_dt_le_1e-5"] >= 0.999, m
    assert m["max_abs_ds"] <= 1e-5, m
    del ts, te, ts_r, te_r, ds, rel_t
    torch.cuda.empty_cache()

    # ---- loss and gradients (requires_grad=True path) at 2^18 rays
    n_rays = 1 << 18
    fns, scales, _ = make_fns(n_rays)
    out = {}
    for name, est, mod in (("ours", nf.PropNetEstimator().to(device), nf), ("ref", ref.PropNetEstimator().to(device), ref)):
        t0, t1 = est.sampling(fns, [256, 96], 48, n_rays, near, far, "lindisp", False, True)
        with torch.no_grad():
            sig = 6.0 * torch.exp(-(((t0 + t1) * 0.5 - 2.5) / 0.2) ** 2)
            trans, _ = mod.render_transmittance_from_density(t0, t1, sig)
        loss = est.compute_loss(trans, loss_scaler=1.0)
        grads = torch.autograd.grad(loss, scales)
        out[name] = (float(loss), [float(g) for g in grads])
    (l_o, g_o), (l_r, g_r) = out["ours"], out["ref"]
    m2 = {"loss": [l_o, l_r], "grads": [g_o, g_r],
          "rel_loss": abs(l_o - l_r) / abs(l_r), "rel_grads": [abs(a - b) / max(abs(b), 1e-30) for a, b in zip(g_o, g_r)]}
    record("PropNetEstimator.compute_loss + grads, 2^18 rays, vs the reference's PropNetEstimator", m2)
    # measured: loss 6.4e-8 relative, gradients equal to the last bit
    assert l_r > 0 and m2["rel_loss"] <= 1e-5, m2
    assert all(r <= 1e-5 for r in m2["rel_grads"]), m2
