"""Fused volume-rendering weights, accumulation and pack_info on the GPU against the oracle
and against the reference's own known-answer vectors (reference tests/test_rendering.py,
tests/test_pack.py and the docstrings of nerfacc/volrend.py).

Float tolerance: rtol 1e-5 (north_star) + atol 2e-6 on quantities of magnitude <= 1 — alpha =
1 - exp(-x) carries an absolute error of one ulp of 1.0 whichever exp is used, so a purely
relative bound is not meaningful for tiny alphas.
"""
import numpy as np
import pytest
import torch

import oracle
from tests._util import (assert_as_close_as_reference, assert_close, packed_info_from_counts, ragged_counts,
                         ray_indices_from_counts)

pytestmark = pytest.mark.gpu
RTOL, ATOL = 1e-5, 2e-6


def _scene(n_rays=3000, seed=0, long_rays=((5, 9000), (6, 4096))):
    counts = ragged_counts(n_rays, seed=seed, long_rays=long_rays)
    pi = packed_info_from_counts(counts)
    ri = ray_indices_from_counts(counts)
    n = ri.shape[0]
    rng = np.random.default_rng(seed + 7)
    ts = rng.random(n).astype(np.float32)
    te = (ts + 0.01 + 0.05 * rng.random(n)).astype(np.float32)
    sig = (20.0 * rng.random(n) ** 3).astype(np.float32)
    return counts, pi, ri, ts, te, sig


def _grad64(ri, pi, gw, gT, ga, ts, te, sig, alphas, prefix):
    """d/d(sigmas | alphas) of sum(w gw + T gT [+ alpha ga]) in float64 (torch autograd on the CPU;
    segmented exclusive sums as a global cumsum minus the ray's offset)."""
    ri_t, pi_t = torch.from_numpy(ri), torch.from_numpy(pi)
    f = lambda a: None if a is None else torch.from_numpy(np.asarray(a, np.float64))
    first = pi_t[:, 0].clamp_max(max(ri_t.numel() - 1, 0))
    if alphas is None:
        x = f(sig).requires_grad_(True)
        sd = x * (f(te) - f(ts))
        a = 1.0 - torch.exp(-sd)
        v = sd
    else:
        x = f(alphas).requires_grad_(True)
        a = x
        v = -torch.log1p(-x)
    cs = torch.cumsum(v, 0) - v
    T = torch.exp(-(cs - cs[first][ri_t]))
    if prefix is not None:
        T = T * f(prefix)
    w = T * a
    loss = (w * f(gw)).sum() + (T * f(gT)).sum() + ((a * f(ga)).sum() if ga is not None else 0.0)
    (g,) = torch.autograd.grad(loss, x)
    return g.numpy()


def _t(a, device, grad=False):
    t = torch.from_numpy(np.ascontiguousarray(a)).to(device)
    if grad:
        t.requires_grad_(True)
    return t


@pytest.mark.parametrize("with_prefix", [False, True])
@pytest.mark.parametrize("api", ["packed_info", "ray_indices"])
def test_weight_from_density_fwd_bwd(device, with_prefix, api):
    import nerfacc_b200 as nf

    counts, pi, ri, ts, te, sig = _scene()
    n = ri.shape[0]
    rng = np.random.default_rng(3)
    prefix = rng.random(n).astype(np.float32) if with_prefix else None
    gw, gT, ga = (rng.standard_normal(n).astype(np.float32) for _ in range(3))
    s = _t(sig, device, grad=True)
    kw = dict(packed_info=_t(pi, device)) if api == "packed_info" else dict(
        ray_indices=_t(ri, device), n_rays=len(counts))
    if api == "packed_info":
        from nerfacc_b200._packing import mark_canonical
        mark_canonical(kw["packed_info"])
    w, T, a = nf.render_weight_from_density(_t(ts, device), _t(te, device), s,
                                            prefix_trans=None if prefix is None else _t(prefix, device), **kw)
    w0, T0, a0 = oracle.render_weight_from_density(ts, te, sig, pi, prefix)
    assert_close(w.detach().cpu().numpy(), w0, RTOL, ATOL, "weights")
    assert_close(T.detach().cpu().numpy(), T0, RTOL, ATOL, "trans")
    assert_close(a.detach().cpu().numpy(), a0, RTOL, ATOL, "alphas")
    (w * _t(gw, device) + T * _t(gT, device) + a * _t(ga, device)).sum().backward()
    gs0 = oracle.render_weight_from_density_bwd(gw, gT, ga, ts, te, sig, pi, prefix)
    # the gradient subtracts a reverse sum from a direct term: any fp32 evaluation carries the
    # cancellation; ours must be as close to float64 as the oracle's sequential fp32 is
    gs64 = _grad64(ri, pi, gw, gT, ga, ts, te, sig, None, prefix)
    assert_as_close_as_reference(s.grad.cpu().numpy(), gs0, gs64, "grad sigmas")


@pytest.mark.parametrize("with_prefix", [False, True])
def test_weight_from_alpha_fwd_bwd(device, with_prefix):
    import nerfacc_b200 as nf
    from nerfacc_b200._packing import mark_canonical

    counts, pi, ri, ts, te, sig = _scene(seed=1)
    n = ri.shape[0]
    rng = np.random.default_rng(4)
    alphas = (rng.random(n) ** 4).astype(np.float32)
    prefix = rng.random(n).astype(np.float32) if with_prefix else None
    gw, gT = (rng.standard_normal(n).astype(np.float32) for _ in range(2))
    a = _t(alphas, device, grad=True)
    pit = mark_canonical(_t(pi, device))
    w, T = nf.render_weight_from_alpha(a, packed_info=pit,
                                       prefix_trans=None if prefix is None else _t(prefix, device))
    w0, T0 = oracle.render_weight_from_alpha(alphas, pi, prefix)
    assert_close(w.detach().cpu().numpy(), w0, RTOL, ATOL, "weights")
    assert_close(T.detach().cpu().numpy(), T0, RTOL, ATOL, "trans")
    (w * _t(gw, device) + T * _t(gT, device)).sum().backward()
    ga0 = oracle.render_weight_from_alpha_bwd(gw, gT, alphas, pi, prefix)
    ga64 = _grad64(ri, pi, gw, gT, None, None, None, None, alphas, prefix)
    assert_as_close_as_reference(a.grad.cpu().numpy(), ga0, ga64, "grad alphas")


def test_noncanonical_packed_info_falls_back(device):
    """A hand-made packed_info (not registered as canonical) takes the general kernels."""
    import nerfacc_b200 as nf

    counts, pi, ri, ts, te, sig = _scene(n_rays=200, long_rays=())
    w, T, a = nf.render_weight_from_density(_t(ts, device), _t(te, device), _t(sig, device),
                                            packed_info=_t(pi, device))
    w0, T0, a0 = oracle.render_weight_from_density(ts, te, sig, pi)
    assert_close(w.cpu().numpy(), w0, RTOL, ATOL, "weights")


def test_visibility_kats(device):
    """Reference tests/test_rendering.py:8-34 and docstring volrend.py:403-409."""
    import nerfacc_b200 as nf

    ri = torch.tensor([0, 2, 2, 2, 2], dtype=torch.int64, device=device)
    alphas = torch.tensor([0.4, 0.3, 0.8, 0.8, 0.5], dtype=torch.float32, device=device)
    vis = nf.render_visibility_from_alpha(alphas, ray_indices=ri, early_stop_eps=0.03, alpha_thre=0.0)
    assert vis.tolist() == [True, True, True, True, False]
    vis = nf.render_visibility_from_alpha(alphas, ray_indices=ri, early_stop_eps=0.05, alpha_thre=0.35)
    assert vis.tolist() == [True, False, True, True, False]
    alphas = torch.tensor([0.4, 0.8, 0.1, 0.8, 0.1, 0.0, 0.9], device=device)
    ri = torch.tensor([0, 0, 0, 1, 1, 2, 2], device=device)
    vis = nf.render_visibility_from_alpha(alphas, ray_indices=ri, early_stop_eps=0.3, alpha_thre=0.2)
    assert vis.tolist() == [True, True, False, True, False, False, True]
    ts = torch.arange(7, dtype=torch.float32, device=device)
    vis = nf.render_visibility_from_density(ts, ts + 1, alphas, ray_indices=ri, early_stop_eps=0.3, alpha_thre=0.2)
    assert vis.tolist() == [True, True, False, True, False, False, True]


def test_visibility_large(device):
    import nerfacc_b200 as nf

    counts, pi, ri, ts, te, sig = _scene(seed=2)
    vis = nf.render_visibility_from_density(_t(ts, device), _t(te, device), _t(sig, device),
                                            ray_indices=_t(ri, device), n_rays=len(counts),
                                            early_stop_eps=1e-2, alpha_thre=1e-2).cpu().numpy()
    w0, T0, a0 = oracle.render_weight_from_density(ts, te, sig, pi)
    ref = oracle.visibility(T0, a0, 1e-2, 1e-2)
    # samples whose T or alpha sits within float noise of a threshold may legitimately flip
    safe = (np.abs(T0 - 1e-2) > 1e-6) & (np.abs(a0 - 1e-2) > 1e-6)
    assert (vis[safe] == ref[safe]).all()
    assert (~safe).sum() < 50


def test_weight_kats_and_grads(device):
    """Reference tests/test_rendering.py:38-57, 110-193 (six API variants, gradient KAT)."""
    import nerfacc_b200 as nf

    ri = torch.tensor([0, 2, 2, 2, 2], dtype=torch.int64, device=device)
    pinfo = torch.tensor([[0, 1], [1, 0], [1, 4]], dtype=torch.long, device=device)
    alphas = torch.tensor([0.4, 0.3, 0.8, 0.8, 0.5], dtype=torch.float32, device=device)
    w, _ = nf.render_weight_from_alpha(alphas, ray_indices=ri, n_rays=3)
    tgt = torch.tensor([1.0 * 0.4, 1.0 * 0.3, 0.7 * 0.8, 0.14 * 0.8, 0.028 * 0.5], device=device)
    assert torch.allclose(w, tgt)

    sigmas = torch.tensor([0.4, 0.8, 0.1, 0.8, 0.1], device=device, requires_grad=True)
    t_starts = torch.rand_like(sigmas)
    t_ends = t_starts + 1.0
    w_ref = torch.tensor([0.3297, 0.5507, 0.0428, 0.2239, 0.0174], device=device)
    g_ref = torch.tensor([0.6703, 0.1653, 0.1653, 0.1653, 0.1653], device=device)

    def check(weights):
        weights.sum().backward()
        g = sigmas.grad.clone()
        sigmas.grad.zero_()
        assert torch.allclose(w_ref, weights, atol=1e-4)
        assert torch.allclose(g_ref, g, atol=1e-4)

    for kw in (dict(ray_indices=ri, n_rays=3), dict(packed_info=pinfo, n_rays=3)):
        trans, _ = nf.render_transmittance_from_density(t_starts, t_ends, sigmas, **kw)
        check(trans * (1.0 - torch.exp(-sigmas * (t_ends - t_starts))))
        check(nf.render_weight_from_density(t_starts, t_ends, sigmas, **kw)[0])
        a = 1.0 - torch.exp(-sigmas * (t_ends - t_starts))
        check(nf.render_weight_from_alpha(a, **kw)[0])


@pytest.mark.parametrize("dim", [None, 1, 2, 3, 4, 7])
@pytest.mark.parametrize("known_layout", [True, False])
def test_accumulate_fwd_bwd(device, dim, known_layout):
    import nerfacc_b200 as nf

    counts, pi, ri, ts, te, sig = _scene(seed=3)
    n, R = ri.shape[0], len(counts)
    rng = np.random.default_rng(9)
    wts = rng.random(n).astype(np.float32)
    vals = None if dim is None else rng.standard_normal((n, dim)).astype(np.float32)
    w = _t(wts, device, grad=True)
    v = None if vals is None else _t(vals, device, grad=True)
    rit = _t(ri, device)
    if known_layout:
        assert (nf.pack_info(rit, R).cpu().numpy() == pi).all()  # also registers the layout
    out = nf.accumulate_along_rays(w, v, rit, R)
    ref = oracle.accumulate_along_rays(wts, vals, ri, R)
    scale = float(np.abs(ref).max())
    assert_close(out.detach().cpu().numpy(), ref, 1e-5, 2e-6 * scale, "accumulate")
    go = rng.standard_normal(ref.shape).astype(np.float32)
    (out * _t(go, device)).sum().backward()
    gw0, gv0 = oracle.accumulate_along_rays_bwd(go, wts, vals, ri)
    assert_close(w.grad.cpu().numpy(), gw0, 1e-5, 1e-5, "g_weights")
    if vals is not None:
        assert_close(v.grad.cpu().numpy(), gv0, 1e-5, 1e-6, "g_values")
    # in-place variant on the same layout: outputs += ray sums (rows of empty rays keep their value)
    acc = torch.full((R, ref.shape[1]), 0.25, device=device)
    nf.accumulate_along_rays_(w.detach(), None if v is None else v.detach(), rit, acc)
    # (rows wider than 4 go through atomics: the summation order, and with it the last bits, vary from run to run)
    assert_close(acc.cpu().numpy(), ref + 0.25, 1e-5, (4e-6 if (dim or 1) > 4 else 2e-6) * max(scale, 1.0), "accumulate_")
    assert (acc.cpu().numpy()[counts == 0] == 0.25).all()


def test_accumulate_kat_and_inplace(device):
    """Reference tests/test_rendering.py:87-106 (incl. an empty ray) + the in-place variant."""
    import nerfacc_b200 as nf

    ri = torch.tensor([0, 2, 2, 2, 2], dtype=torch.int64, device=device)
    w = torch.tensor([0.4, 0.3, 0.8, 0.8, 0.5], dtype=torch.float32, device=device)
    v = torch.rand((5, 2), device=device)
    out = nf.accumulate_along_rays(w, values=v, ray_indices=ri, n_rays=3)
    assert out.shape == (3, 2)
    assert torch.allclose(out[0], w[0, None] * v[0])
    assert (out[1] == 0).all()
    assert torch.allclose(out[2], (w[1:, None] * v[1:]).sum(dim=0))
    acc = torch.ones((3, 2), device=device)
    nf.accumulate_along_rays_(w, values=v, ray_indices=ri, outputs=acc)
    assert torch.allclose(acc, out + 1.0)
    unsorted = torch.tensor([2, 0, 2, 1, 2], dtype=torch.int64, device=device)
    out = nf.accumulate_along_rays(w, values=v, ray_indices=unsorted, n_rays=3)
    ref = torch.zeros((3, 2), device=device).index_add_(0, unsorted, w[:, None] * v)
    assert torch.allclose(out, ref)


def test_pack_info(device):
    """Reference tests/test_pack.py:8-18 and docstring pack.py:29-32, plus a ragged case."""
    import nerfacc_b200 as nf

    ri = torch.tensor([0, 2, 2, 2, 2], dtype=torch.int64, device=device)
    assert nf.pack_info(ri, n_rays=3).tolist() == [[0, 1], [1, 0], [1, 4]]
    ri = torch.tensor([0, 0, 1, 1, 1, 2, 2, 2, 2], device=device)
    assert nf.pack_info(ri, n_rays=3).tolist() == [[0, 2], [2, 3], [5, 4]]
    assert nf.pack_info(ri).tolist() == [[0, 2], [2, 3], [5, 4]]
    counts = ragged_counts(70000, seed=11, long_rays=[(100, 20000)])
    pi = packed_info_from_counts(counts)
    got = nf.pack_info(_t(ray_indices_from_counts(counts), device), len(counts)).cpu().numpy()
    assert (got == pi).all()
    empty = nf.pack_info(torch.zeros(0, dtype=torch.long, device=device), 5)
    assert empty.tolist() == [[0, 0]] * 5


def _pack_info_torch(ri, n_rays):
    """The reference's formula (pack.py:43-46) in torch."""
    cnt = torch.zeros((n_rays,), dtype=torch.long, device=ri.device)
    cnt.index_add_(0, ri, torch.ones_like(ri))
    return torch.stack([cnt.cumsum(0) - cnt, cnt], -1)


@pytest.mark.parametrize("case", ["sorted", "gaps", "tail", "unsorted", "one_decrease", "single_ray"])
def test_pack_info_sorted_pass_and_fallback(device, case):
    """Sorted indices take the two plain passes; an index that decreases anywhere is noticed on the device
    and the general pass redoes the work: either way the result is the reference's formula."""
    import nerfacc_b200 as nf

    g = torch.Generator().manual_seed(5)
    n_rays = 50000
    if case == "sorted":
        ri = torch.sort(torch.randint(0, n_rays, (400003,), generator=g)).values
    elif case == "gaps":  # a few far-apart rays: long stretches of empty rays before, between and after
        ri = torch.sort(torch.tensor([7, 7, 7, 123, 30000, 30000, 30001, 49000])).values
    elif case == "tail":  # nothing beyond ray 10: 49989 trailing empty rays; leading empty rays too
        ri = torch.sort(torch.randint(5, 11, (1000,), generator=g)).values
    elif case == "unsorted":
        ri = torch.randint(0, n_rays, (400003,), generator=g)
    elif case == "one_decrease":  # sorted except for the very last element
        ri = torch.sort(torch.randint(0, n_rays, (100000,), generator=g)).values
        ri[-1] = 3
    else:
        ri = torch.full((70000,), 42, dtype=torch.long)
    ri = ri.to(device)
    got = nf.pack_info(ri, n_rays)
    assert torch.equal(got, _pack_info_torch(ri, n_rays))
    # and again on the same workspace (the flag is reset every call): sorted after unsorted, unsorted after sorted
    other = torch.sort(ri).values if case in ("unsorted", "one_decrease") else ri.flip(0).contiguous()
    assert torch.equal(nf.pack_info(other, n_rays), _pack_info_torch(other, n_rays))
    assert torch.equal(nf.pack_info(ri[:], n_rays), got)


def test_rendering_end_to_end(device):
    import nerfacc_b200 as nf

    counts, pi, ri, ts, te, sig = _scene(seed=5)
    R = len(counts)
    rng = np.random.default_rng(12)
    rgb = rng.random((ri.shape[0], 3)).astype(np.float32)
    s = _t(sig, device, grad=True)
    c = _t(rgb, device, grad=True)
    bk = torch.tensor([1.0, 0.5, 0.25], device=device)
    colors, opac, depth, extras = nf.rendering(_t(ts, device), _t(te, device), _t(ri, device), R,
                                               rgb_sigma_fn=lambda a, b, r: (c, s), render_bkgd=bk)
    w0, T0, a0 = oracle.render_weight_from_density(ts, te, sig, pi)
    col0 = oracle.accumulate_along_rays(w0, rgb, ri, R)
    op0 = oracle.accumulate_along_rays(w0, None, ri, R)
    dp0 = oracle.accumulate_along_rays(w0, ((ts + te) / 2.0)[:, None], ri, R)
    assert_close(opac.detach().cpu().numpy(), op0, 1e-5, 2e-6, "opacity")
    assert_close(colors.detach().cpu().numpy(), col0 + np.array([1.0, 0.5, 0.25], np.float32) * (1 - op0),
                 1e-5, 4e-6, "colors")
    assert_close(depth.detach().cpu().numpy(), dp0 / np.maximum(op0, np.finfo(np.float32).eps), 2e-5, 4e-6, "depth")
    assert set(extras) == {"weights", "alphas", "trans", "sigmas", "rgbs"}
    (colors.sum() + opac.sum()).backward()
    assert torch.isfinite(s.grad).all() and torch.isfinite(c.grad).all()


@pytest.mark.parametrize("mode", ["density", "alpha"])
@pytest.mark.parametrize("path", ["bits", "bits_odd", "per_ray", "per_ray_L4", "per_ray_L32", "tiles"])
def test_rendering_fused_composite_vs_oracle(device, mode, path, monkeypatch):
    """rendering() on flattened inputs — the warp-tile kernels (csrc/render_bits.cu, the default
    for dense layouts), the per-ray kernels (csrc/render_rays.cu, lanes per ray chosen / forced)
    and the flat tile kernels (volrend.cu + composite.cu) — against the oracle's composition of
    volrend.py:140-156, incl. empty rays, a long ray that spans several tiles, a background
    colour, and an extra loss on extras['weights']."""
    import nerfacc_b200 as nf
    from nerfacc_b200 import _render_ops

    if path.startswith("bits"):
        monkeypatch.setattr(_render_ops, "_RENDER_MODE", "bits")
    elif path == "tiles":
        monkeypatch.setattr(_render_ops, "_RENDER_MODE", "flat")
    else:
        monkeypatch.setattr(_render_ops, "_RENDER_MODE", "rays")
        if path != "per_ray":
            monkeypatch.setattr(_render_ops, "_RAYS_L", int(path.split("L")[1]))

    if path == "bits_odd":
        # total not a multiple of 4, rays ending exactly on / one past warp-tile and overhang
        # boundaries, runs of 1-sample rays, rays of 128..700 samples (look-back between warps)
        counts = ragged_counts(2000, mean=9, seed=11, long_rays=((3, 509), (4, 3), (5, 128), (6, 127), (7, 700),
                                                                 (900, 2049), (901, 1), (902, 1), (903, 1),
                                                                 (1999, 131)))
        if counts.sum() % 4 == 0:
            counts[10] += 1
    else:
        counts = ragged_counts(3000, mean=14, seed=5, long_rays=((17, 9000), (1500, 1)))
    pi = packed_info_from_counts(counts)
    ri = ray_indices_from_counts(counts)
    n, R = int(counts.sum()), len(counts)
    rng = np.random.default_rng(9)
    ts = np.sort(rng.random(n).astype(np.float32) * 4)
    te = (ts + rng.random(n).astype(np.float32) * 0.02 + 1e-3).astype(np.float32)
    field = (rng.random(n) * (8.0 if mode == "density" else 0.3)).astype(np.float32)
    rgb = rng.random((n, 3)).astype(np.float32)
    gc, go, gd = (rng.random((R, k)).astype(np.float32) for k in (3, 1, 1))
    gwx = rng.random(n).astype(np.float32)
    bk = np.array([0.1, 0.5, 0.9], np.float32)

    t = lambda a: torch.from_numpy(a).to(device)
    f_t, rgb_t = t(field).requires_grad_(True), t(rgb).requires_grad_(True)
    fn = (lambda a, b, r: (rgb_t, f_t))
    kw = dict(rgb_sigma_fn=fn) if mode == "density" else dict(rgb_alpha_fn=fn)
    colors, opac, depth, extras = nf.rendering(t(ts), t(te), t(ri), R, render_bkgd=t(bk), **kw)
    loss = (colors * t(gc)).sum() + (opac * t(go)).sum() + (depth * t(gd)).sum() + (extras["weights"] * t(gwx)).sum()
    loss.backward()

    if mode == "density":
        w, T, a = oracle.render_weight_from_density(ts, te, field, pi)
    else:
        w, T = oracle.render_weight_from_alpha(field, pi)
    tmid = ((ts + te) / 2)[:, None].astype(np.float32)
    c0 = oracle.accumulate_along_rays(w, rgb, ri, R)
    o0 = oracle.accumulate_along_rays(w, None, ri, R)
    d0 = oracle.accumulate_along_rays(w, tmid, ri, R)
    eps = np.finfo(np.float32).eps
    den = np.maximum(o0, eps)
    assert_close(extras["weights"].detach().cpu().numpy(), w, 1e-5, 1e-7, "weights")
    assert_close(colors.detach().cpu().numpy(), c0 + bk * (1 - o0), 2e-5, 1e-5, "colors")
    assert_close(opac.detach().cpu().numpy(), o0, 2e-5, 1e-5, "opacity")
    assert_close(depth.detach().cpu().numpy(), d0 / den, 1e-4, 1e-5, "depth")
    assert (opac.detach().cpu().numpy()[counts == 0] == 0).all()
    # backward: colors = c0 + bk*(1-o0), depth = d0/den
    g_c0 = gc
    g_o0 = go - (gc * bk).sum(-1, keepdims=True) - np.where(o0 > eps, gd * d0 / (den * den), 0.0)
    g_d0 = gd / den
    gw1, g_rgb = oracle.accumulate_along_rays_bwd(g_c0, w, rgb, ri)
    gw2, _ = oracle.accumulate_along_rays_bwd(g_o0.astype(np.float32), w, None, ri)
    gw3, _ = oracle.accumulate_along_rays_bwd(g_d0.astype(np.float32), w, tmid, ri)
    gw = (gw1 + gw2 + gw3 + gwx).astype(np.float32)
    if mode == "density":
        g_f = oracle.render_weight_from_density_bwd(gw, None, None, ts, te, field, pi)
    else:
        g_f = oracle.render_weight_from_alpha_bwd(gw, None, field, pi)
    assert_close(rgb_t.grad.cpu().numpy(), g_rgb, 1e-5, 1e-7, "g_rgb")
    scale = float(np.abs(g_f).max())
    # (bits_odd has 1-2 sample rays of tiny opacity: depth = d0 / o0 makes the per-ray gradient rows
    # themselves — torch on the GPU vs numpy here — differ by a few 1e-4 relative after cancellation)
    assert_close(f_t.grad.cpu().numpy(), g_f, 1e-4, (2e-5 if path == "bits_odd" else 2e-6) * scale, "g_field")


def test_depth_normalize_matches_torch(device):
    """nfb_depth_normalize_fwd / _bwd against torch's own depth / opacity.clamp_min(eps) and its
    autograd (volrend.py:147-150), incl. rays of zero and sub-eps opacity."""
    from nerfacc_b200 import _C

    R = 10007
    g = torch.Generator(device=device); g.manual_seed(3)
    op = torch.rand(R, 1, device=device, generator=g)
    op[::13] = 0.0
    op[5::17] = 3e-8
    d0 = torch.rand(R, 1, device=device, generator=g) * 4
    g_op = torch.randn(R, 1, device=device, generator=g)
    g_dp = torch.randn(R, 1, device=device, generator=g)
    eps = torch.finfo(torch.float32).eps
    o_t, d_t = op.clone().requires_grad_(True), d0.clone().requires_grad_(True)
    ref = d_t / o_t.clamp_min(eps)
    torch.autograd.backward([ref, o_t * 1.0], [g_dp, g_op])
    depth = d0.clone()
    lib = _C.lib()
    with torch.cuda.device(device):
        _C.check(lib.nfb_depth_normalize_fwd(op.data_ptr(), depth.data_ptr(), R, _C.stream_ptr(device)), "fwd")
        g_o0, g_d0 = torch.empty_like(op), torch.empty_like(op)
        _C.check(lib.nfb_depth_normalize_bwd(g_op.data_ptr(), g_dp.data_ptr(), op.data_ptr(), depth.data_ptr(), R,
                                             g_o0.data_ptr(), g_d0.data_ptr(), _C.stream_ptr(device)), "bwd")
    assert torch.equal(depth, ref.detach())
    assert_close(g_d0.cpu().numpy(), d_t.grad.cpu().numpy(), 1e-6, 0.0, "g_depth_sum")
    scale = float(o_t.grad.abs().max())
    assert_close(g_o0.cpu().numpy(), o_t.grad.cpu().numpy(), 1e-5, 1e-7 * scale, "g_opacity_sum")
