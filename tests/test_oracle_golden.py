"""Pin the CPU oracle (oracle/) before trusting it as the parity checker.

Three sources, all of them the reference's own:
  (a) known-answer vectors of the reference's tests and docstrings (file:line cited per test),
  (b) tests/golden/golden_cpu.npz — the UNMODIFIED Python reference run on CPU
      (tests/golden/make_golden_cpu.py),
  (c) tests/golden/golden_gpu.npz — the reference's own CUDA kernels run on a B200
      (tests/golden/make_golden_gpu.py through oracle/_ref/nerfacc_cuda.so), bit-exact for the
      traversal.
Runs without a GPU.
"""
import os

import numpy as np
import pytest

import oracle
from tests._util import packed_info_from_counts, ray_indices_from_counts

HERE = os.path.dirname(os.path.abspath(__file__))
RTOL, ATOL = 1e-5, 1e-6  # north_star: 1e-5 relative in fp32


@pytest.fixture(scope="module")
def G():
    return np.load(os.path.join(HERE, "golden", "golden_cpu.npz"))


def _uniform_info(R, N):
    return packed_info_from_counts(np.full(R, N, np.int64))


# ----------------------------------------------------------------------------- (a) KATs
def test_kat_pack_info():
    """reference tests/test_pack.py:8-18 and nerfacc/pack.py:29-32."""
    assert (oracle.pack_info([0, 2, 2, 2, 2], 3) == [[0, 1], [1, 0], [1, 4]]).all()
    assert (oracle.pack_info([0, 0, 1, 1, 1, 2, 2, 2, 2], 3) == [[0, 2], [2, 3], [5, 4]]).all()


def test_kat_scans_docstrings():
    """nerfacc/scan.py:36-39, 78-81, 127-130, 170-173."""
    x = np.arange(1, 10, dtype=np.float32)
    pi = np.array([[0, 2], [2, 3], [5, 4]])
    assert (oracle.inclusive_sum(x, pi) == [1, 3, 3, 7, 12, 6, 13, 21, 30]).all()
    assert (oracle.exclusive_sum(x, pi) == [0, 1, 0, 3, 7, 0, 6, 13, 21]).all()
    assert (oracle.inclusive_prod(x, pi) == [1, 2, 3, 12, 60, 6, 42, 336, 3024]).all()
    assert (oracle.exclusive_prod(x, pi) == [1, 1, 1, 3, 12, 1, 6, 42, 336]).all()


def test_kat_visibility_and_weights():
    """reference tests/test_rendering.py:8-57 and volrend.py docstrings :192-195, 298-302."""
    a = np.array([0.4, 0.3, 0.8, 0.8, 0.5], np.float32)
    pi = oracle.pack_info([0, 2, 2, 2, 2], 3)
    w, T = oracle.render_weight_from_alpha(a, pi)
    np.testing.assert_allclose(T, [1.0, 1.0, 0.7, 0.14, 0.028], rtol=1e-5)
    np.testing.assert_allclose(w, [0.4, 0.3, 0.56, 0.112, 0.014], rtol=1e-5)
    assert (oracle.visibility(T, a, 0.03, 0.0) == [True, True, True, True, False]).all()
    assert (oracle.visibility(T, a, 0.05, 0.35) == [True, False, True, True, False]).all()
    a2 = np.array([0.4, 0.8, 0.1, 0.8, 0.1, 0.0, 0.9], np.float32)
    pi2 = oracle.pack_info([0, 0, 0, 1, 1, 2, 2], 3)
    w2, T2 = oracle.render_weight_from_alpha(a2, pi2)
    np.testing.assert_allclose(T2, [1.0, 0.6, 0.12, 1.0, 0.2, 1.0, 1.0], rtol=1e-5)
    np.testing.assert_allclose(w2, [0.4, 0.48, 0.012, 0.8, 0.02, 0.0, 0.9], rtol=1e-5)
    assert (oracle.visibility(T2, a2, 0.3, 0.2) == [True, True, False, True, False, False, True]).all()


def test_kat_gradients():
    """reference tests/test_rendering.py:110-193: weights and d(sum weights)/d sigmas."""
    sig = np.array([0.4, 0.8, 0.1, 0.8, 0.1], np.float32)
    ts = np.array([0.1, 0.3, 0.5, 0.2, 0.9], np.float32)
    te = ts + 1.0
    pi = np.array([[0, 1], [1, 0], [1, 4]])
    w, T, a = oracle.render_weight_from_density(ts, te, sig, pi)
    np.testing.assert_allclose(w, [0.3297, 0.5507, 0.0428, 0.2239, 0.0174], atol=1e-4)
    g = oracle.render_weight_from_density_bwd(np.ones(5, np.float32), None, None, ts, te, sig, pi)
    np.testing.assert_allclose(g, [0.6703, 0.1653, 0.1653, 0.1653, 0.1653], atol=1e-4)
    # same through the alpha path chained with alpha = 1 - exp(-sigma * delta)
    ga = oracle.render_weight_from_alpha_bwd(np.ones(5, np.float32), None, a, pi)
    g2 = ga * np.exp(-sig * (te - ts)) * (te - ts)
    np.testing.assert_allclose(g2, [0.6703, 0.1653, 0.1653, 0.1653, 0.1653], atol=1e-4)


def test_kat_accumulate_with_empty_ray():
    """reference tests/test_rendering.py:87-106."""
    w = np.array([0.4, 0.3, 0.8, 0.8, 0.5], np.float32)
    v = np.random.default_rng(0).random((5, 2)).astype(np.float32)
    out = oracle.accumulate_along_rays(w, v, [0, 2, 2, 2, 2], 3)
    np.testing.assert_allclose(out[0], w[0] * v[0], rtol=1e-6)
    assert (out[1] == 0).all()
    np.testing.assert_allclose(out[2], (w[1:, None] * v[1:]).sum(0), rtol=1e-6)


def test_kat_searchsorted_and_importance_sampling_docstrings():
    """nerfacc/pdf.py:40-56 (packed searchsorted) and :108-120 (importance_sampling)."""
    key = np.array([0.0, 1.0, 0.0, 1.0, 2.0], np.float32)
    key_info = np.array([[0, 2], [2, 3]])
    q = np.array([0.5, 1.5, 2.5], np.float32)
    q_info = np.array([[0, 1], [1, 2]])
    il, ir = oracle.searchsorted(key, q, key_info, q_info)
    assert (il == [0, 3, 3]).all() and (ir == [1, 4, 4]).all()
    cdfs = np.array([0.0, 0.5, 0.0, 0.5, 1.0], np.float32)
    edges, cen = oracle.importance_sampling(key, cdfs, 2, packed_info=key_info)
    np.testing.assert_allclose(edges, [[0.0, 0.5, 1.0], [0.0, 1.0, 2.0]], atol=1e-6)
    np.testing.assert_allclose(cen, [[0.25, 0.75], [0.5, 1.5]], atol=1e-6)


def test_kat_near_far_traversal():
    """reference tests/test_grid.py:135-159: 1x1x1 grid, edges stay within [near, far]."""
    o = np.array([[-1.0, 0.0, 0.0]], np.float32)
    d = np.array([[1.0, 0.01, 0.01]], np.float32)
    d /= np.linalg.norm(d)
    r = oracle.traverse_grids(o, d, np.ones((1, 1, 1, 1), bool), np.array([[0, 0, 0, 1, 1, 1]], np.float32),
                              np.array([1.2], np.float32), np.array([1.5], np.float32), step_size=0.05)
    v = r["intervals"]["vals"]
    assert v.size > 0 and (v >= 1.2 - 0.025).all() and (v <= 1.5 + 0.025).all()
    assert abs(v[0] - 1.2) < 1e-6 and abs(v[-1] - 1.4999998) < 1e-5  # SURVEY §8c evidence


def test_property_traverse_samples_land_in_occupied_cells():
    """reference tests/test_grid.py:39-68: every sample midpoint of a march through 4 nested
    32^3 random grids lies in an occupied cell of the finest level that contains it."""
    rng = np.random.default_rng(42)
    n, m, res = 10, 4, 32
    o = rng.standard_normal((n, 3)).astype(np.float32)
    d = rng.standard_normal((n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=-1, keepdims=True)
    bins = rng.random((m, res, res, res)) < 0.5
    aabbs = np.stack([np.array([-1, -1, -1, 1, 1, 1], np.float32) * 2**i for i in range(m)])
    r = oracle.traverse_grids(o, d, bins, aabbs, step_size=1e-3)
    sm = r["samples"]
    assert sm["vals"].size > 10000
    x = o[sm["ray_indices"]] + d[sm["ray_indices"]] * sm["vals"][:, None]
    reach = np.abs(x).max(-1)
    mip = np.clip(np.ceil(np.log2(np.maximum(reach, 1e-9))), 0, None).astype(int)
    ok = mip < m
    # interior points only: skip samples within 1e-4 of a level boundary (float edge effects)
    near_edge = np.abs(reach - 2.0 ** np.round(np.log2(np.maximum(reach, 1e-9)))) < 1e-4
    mip = np.clip(mip, 0, m - 1)
    u = (x / (2.0**mip)[:, None] + 1.0) / 2.0
    ijk = np.clip((u * res).astype(int), 0, res - 1)
    occ = bins[mip, ijk[:, 0], ijk[:, 1], ijk[:, 2]]
    assert occ[ok & ~near_edge].all()
    # intervals/samples consistency: E = S + number of runs, mids are edge midpoints
    iv = r["intervals"]
    ts, te = iv["vals"][iv["is_left"]], iv["vals"][iv["is_right"]]
    assert ts.shape == te.shape == sm["vals"].shape
    assert (sm["vals"] == (te + ts) * np.float32(0.5)).all()


# ----------------------------------------------------------------------------- (b) CPU reference
def test_golden_batched_scans(G):
    x, g = G["scan/x"], G["scan/g"]
    R, N = x.shape
    pi = _uniform_info(R, N)
    for name, op, excl in (("inclusive_sum", "sum", False), ("exclusive_sum", "sum", True),
                           ("inclusive_prod", "prod", False), ("exclusive_prod", "prod", True)):
        y = oracle.scan(x.ravel(), pi, op, excl).reshape(R, N)
        np.testing.assert_allclose(y, G[f"scan/{name}"], rtol=RTOL, atol=ATOL)
        if op == "sum":
            gx = oracle.scan(g.ravel(), pi, "sum", excl, reverse=True).reshape(R, N)
        else:
            gx = oracle.prod_scan_backward(g.ravel(), x.ravel(), y.ravel(), pi, excl).reshape(R, N)
        np.testing.assert_allclose(gx, G[f"scan/{name}_grad"], rtol=1e-4, atol=1e-5)


def test_golden_volrend(G):
    ts, te, sig = G["vr/ts"], G["vr/te"], G["vr/sig"]
    gw, gT, ga = G["vr/gw"], G["vr/gT"], G["vr/ga"]
    R, N = sig.shape
    pi = _uniform_info(R, N)
    f = lambda a: np.ascontiguousarray(a).ravel()
    for tag, pt in (("", None), ("_prefix", np.broadcast_to(G["vr/prefix"], (R, N)))):
        w, T, a = oracle.render_weight_from_density(f(ts), f(te), f(sig), pi, None if pt is None else f(pt))
        np.testing.assert_allclose(w.reshape(R, N), G[f"vr/density{tag}_w"], rtol=RTOL, atol=ATOL)
        np.testing.assert_allclose(T.reshape(R, N), G[f"vr/density{tag}_T"], rtol=RTOL, atol=ATOL)
        np.testing.assert_allclose(a.reshape(R, N), G[f"vr/density{tag}_a"], rtol=RTOL, atol=ATOL)
        w2, T2 = oracle.render_weight_from_alpha(f(G["vr/alphas"]), pi, None if pt is None else f(pt))
        np.testing.assert_allclose(w2.reshape(R, N), G[f"vr/alpha{tag}_w"], rtol=RTOL, atol=ATOL)
        np.testing.assert_allclose(T2.reshape(R, N), G[f"vr/alpha{tag}_T"], rtol=RTOL, atol=ATOL)
    gs = oracle.render_weight_from_density_bwd(f(gw), f(gT), f(ga), f(ts), f(te), f(sig), pi)
    np.testing.assert_allclose(gs.reshape(R, N), G["vr/density_grad"], rtol=1e-4, atol=1e-5)
    gs = oracle.render_weight_from_density_bwd(f(gw), None, None, f(ts), f(te), f(sig), pi)
    np.testing.assert_allclose(gs.reshape(R, N), G["vr/density_grad_wonly"], rtol=1e-4, atol=1e-5)
    gal = oracle.render_weight_from_alpha_bwd(f(gw), f(gT), f(G["vr/alphas"]), pi)
    np.testing.assert_allclose(gal.reshape(R, N), G["vr/alpha_grad"], rtol=1e-4, atol=1e-5)
    # visibility masks: exact except where T sits within float noise of the threshold
    w, T = oracle.render_weight_from_alpha(f(G["vr/alphas"]), pi)
    vis = oracle.visibility(T, f(G["vr/alphas"]), 0.05, 0.02).reshape(R, N)
    safe = (np.abs(T - 0.05) > 1e-6).reshape(R, N)
    assert (vis == G["vr/vis_alpha"])[safe].all()
    w, T, a = oracle.render_weight_from_density(f(ts), f(te), f(sig), pi)
    vis = oracle.visibility(T, a, 0.03, 0.1).reshape(R, N)
    safe = ((np.abs(T - 0.03) > 1e-6) & (np.abs(a - 0.1) > 1e-6)).reshape(R, N)
    assert (vis == G["vr/vis_density"])[safe].all()


def test_golden_accumulate_and_rendering(G):
    counts = G["acc/counts"]
    ri = ray_indices_from_counts(counts)
    w, v, go = G["acc/w"], G["acc/v"], G["acc/go"]
    out = oracle.accumulate_along_rays(w, v, ri, len(counts))
    np.testing.assert_allclose(out, G["acc/out"], rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(oracle.accumulate_along_rays(w, None, ri, len(counts)), G["acc/out_none"],
                               rtol=RTOL, atol=ATOL)
    gw, gv = oracle.accumulate_along_rays_bwd(go, w, v, ri)
    np.testing.assert_allclose(gw, G["acc/gw"], rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(gv, G["acc/gv"], rtol=RTOL, atol=ATOL)
    # rendering() batched == oracle composition (volrend.py:140-156)
    ts, te, sig, rgb = G["vr/ts"], G["vr/te"], G["vr/sig"], G["rend/rgb"]
    R, N = sig.shape
    pi = _uniform_info(R, N)
    rix = ray_indices_from_counts(np.full(R, N))
    wts, T, a = oracle.render_weight_from_density(ts.ravel(), te.ravel(), sig.ravel(), pi)
    colors = oracle.accumulate_along_rays(wts, rgb.reshape(-1, 3), rix, R)
    opac = oracle.accumulate_along_rays(wts, None, rix, R)
    depth = oracle.accumulate_along_rays(wts, ((ts + te) / 2).reshape(-1, 1), rix, R)
    depth = depth / np.maximum(opac, np.finfo(np.float32).eps)
    colors = colors + G["rend/bkgd"] * (1.0 - opac)
    np.testing.assert_allclose(colors, G["rend/colors"], rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(opac, G["rend/opac"], rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(depth, G["rend/depth"], rtol=1e-4, atol=1e-5)


def test_golden_ray_aabb(G):
    tmin, tmax, hits = oracle.ray_aabb_intersect(G["aabb/o"], G["aabb/d"], G["aabb/bb"])
    assert (hits == G["aabb/hits"]).all()
    np.testing.assert_allclose(tmin, G["aabb/tmin"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(tmax, G["aabb/tmax"], rtol=1e-5, atol=1e-6)


def test_golden_importance_sampling_and_loss(G):
    """Tolerance 1e-4 absolute, the reference's own (tests/test_pdf.py:93-94, 126-127)."""
    edges, cen = oracle.importance_sampling(G["pdf/bins"], G["pdf/cdfs"], G["pdf/centers"].shape[1])
    np.testing.assert_allclose(cen, G["pdf/centers"], atol=1e-4)
    np.testing.assert_allclose(edges, G["pdf/edges"], atol=1e-4)
    # _pdf_loss (prop_net.py:232-256) restated on top of the oracle's searchsorted
    t_q, w_q, t_k, w_k = G["loss/t_q"], G["loss/w_q"], G["loss/t_k"], G["loss/w_k"]
    c_q = np.concatenate([np.zeros_like(w_q[:, :1]), np.cumsum(w_q, -1)], -1)
    c_k = np.concatenate([np.zeros_like(w_k[:, :1]), np.cumsum(w_k, -1)], -1)
    il, ir = oracle.searchsorted(t_k, t_q)
    w = c_q[:, 1:] - c_q[:, :-1]
    w_outer = np.take_along_axis(c_k, ir[:, 1:], -1) - np.take_along_axis(c_k, il[:, :-1], -1)
    loss = np.clip(w - w_outer, 0, None) ** 2 / (w + 1e-7)
    np.testing.assert_allclose(loss, G["loss/outer"], atol=1e-4)


def test_golden_occupancy_update(G):
    res = [16, 16, 16]
    occs = G["occ/after_mark"].copy()
    for lvl in range(2):
        cells = G[f"occ/cells{lvl}"]
        x = oracle.occ_cell_points(cells, G[f"occ/u{lvl}"], res, G["occ/aabbs"][lvl])
        np.testing.assert_allclose(x, G[f"occ/x{lvl}"], rtol=1e-6, atol=1e-6)
        occs = oracle.occ_ema_update(occs, lvl * 16**3 + cells, G[f"occ/occ{lvl}"].ravel(), 0.95)
    assert (occs == G["occ/after_update"]).all()
    bins, thre = oracle.occ_binarize(occs, 0.01)
    assert (bins == G["occ/binaries"].ravel()).all()


# ----------------------------------------------------------------------------- (c) reference CUDA on B200
GPU_NPZ = os.path.join(HERE, "golden", "golden_gpu.npz")


@pytest.fixture(scope="module")
def GG():
    if not os.path.exists(GPU_NPZ):
        pytest.skip("tests/golden/golden_gpu.npz not generated yet")
    return np.load(GPU_NPZ)


def test_golden_gpu_scans(GG):
    x, pi = GG["scan/x"], GG["scan/packed_info"]
    for op in ("sum", "prod"):
        for excl in (False, True):
            name = f"scan/{'exclusive' if excl else 'inclusive'}_{op}"
            y = oracle.scan(x, pi, op, excl)
            covered = np.zeros(x.shape, bool)
            for s, c in pi:
                covered[s:s + c] = True
            assert (y[covered] == GG[name][covered]).all(), name  # same association -> bit-exact
            if op == "sum":
                gx = oracle.scan(GG["scan/g"], pi, "sum", excl, reverse=True)
            else:
                gx = oracle.prod_scan_backward(GG["scan/g"], x, GG[name], pi, excl)
            np.testing.assert_allclose(gx[covered], GG[name + "_grad"][covered], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("case", ["nested", "cone", "voxel", "limit", "shell"])
def test_golden_gpu_traversal_bit_exact(GG, case):
    p = f"trav/{case}/"
    kw = dict(step_size=float(GG[p + "step_size"]), cone_angle=float(GG[p + "cone_angle"]))
    limit = int(GG[p + "limit"])
    if limit > 0:
        kw["traverse_steps_limit"] = limit
    near = GG[p + "near"] if p + "near" in GG else None
    far = GG[p + "far"] if p + "far" in GG else None
    shape = tuple(int(v) for v in GG[p + "bins_shape"])
    bins = np.unpackbits(GG[p + "bins"])[: int(np.prod(shape))].reshape(shape).astype(bool)
    r = oracle.traverse_grids(GG[p + "o"], GG[p + "d"], bins, GG[p + "aabbs"], near, far, **kw)
    for part, keys in (("intervals", ("vals", "ray_indices", "is_left", "is_right", "packed_info")),
                       ("samples", ("vals", "ray_indices", "is_valid", "packed_info"))):
        for k in keys:
            ref = GG[f"{p}{part}/{k}"]
            got = r[part][k]
            assert got.shape == ref.shape, (case, part, k, got.shape, ref.shape)
            assert (got == ref).all(), (case, part, k, int((got != ref).sum()))
    has = r["samples"]["packed_info"][:, 1] > 0
    assert (r["termination_planes"][has] == GG[p + "termination_planes"][has]).all()


def test_golden_gpu_aabb_pdf(GG):
    tmin, tmax, hits = oracle.ray_aabb_intersect(GG["aabb/o"], GG["aabb/d"], GG["aabb/bb"])
    assert (hits == GG["aabb/hits"]).all()
    assert (tmin == GG["aabb/tmin"]).all() and (tmax == GG["aabb/tmax"]).all()
    edges, cen = oracle.importance_sampling(GG["pdf/vals"], GG["pdf/cdfs"], GG["pdf/centers"].shape[1])
    np.testing.assert_allclose(cen, GG["pdf/centers"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(edges, GG["pdf/edges"], rtol=1e-5, atol=1e-6)
    il, ir = oracle.searchsorted(GG["pdf/vals"], GG["pdf/query"])
    assert (il == GG["pdf/ids_left"]).all() and (ir == GG["pdf/ids_right"]).all()
