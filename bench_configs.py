"""Side legs of bench.py (measurement tooling, not a product path): every BASELINE.json config
other than the headline one, each next to the reference's OWN CUDA kernels on the same tensors
(the unmodified extension in oracle/_ref, the like-for-like bar of SURVEY.md §8d), the pre-flight
parity check of the headline step at its full size, the cfg4 sharded-update check for N > 1, and
the cfg1 CPU baseline (the reference's batched PyTorch path on the host cores).

Nothing here runs inside bench.py's timed region.  Timing method of the `configs` block: CUDA
events on the current stream, 3 warm-ups, then 5 groups of 5 back-to-back calls, median group /
5; every S-sized operand is larger than L2 (126 MB).  Algorithmic bytes as in DESIGN.md §4.
"""
from __future__ import annotations

import math
import os
import statistics
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))


def timeit(fn, warm=3, groups=5, reps=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(groups):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        b.synchronize()
        ts.append(a.elapsed_time(b) / reps)
    return statistics.median(ts)


def native_ms(fn, reps=5):
    """Time spent inside this library's kernels during one call of fn: CUDA events recorded around
    every C-ABI entry point (nerfacc_b200/_C.py prof hooks), summed, averaged over `reps` calls.
    Returns (ms per call, {entry point: launches per call})."""
    from nerfacc_b200 import _C

    torch.cuda.synchronize()
    _C.prof_enable(True)
    try:
        for _ in range(reps):
            fn()
        prof = _C.prof_collect()
    finally:
        _C.prof_enable(False)
    return sum(t for _, t in prof.values()) / reps, {k: c / reps for k, (c, _) in prof.items()}


def _entry(algo_bytes, ours, ref, peak, note=None, **extra):
    """ms: the public call, back to back (host launch cost included: for sub-100 us kernels the
    Python side can be the limit); kernel_ms: the native kernels alone; gbs / frac: algorithmic
    bytes over kernel_ms, the roofline figure; frac_of_call: the same over ms."""
    t_o = timeit(ours)
    t_k, names = native_ms(ours)
    t_r = timeit(ref) if ref is not None else None
    if t_k <= 0.0:
        t_k = t_o
    gbs = algo_bytes / (t_k * 1e-3) / 1e9
    ent = {"ms": t_o, "kernel_ms": t_k, "bytes": int(algo_bytes), "gbs": gbs, "frac": gbs / peak,
           "frac_of_call": algo_bytes / (t_o * 1e-3) / 1e9 / peak, "native_calls": names, "ref_cuda_ms": t_r,
           "speedup_vs_ref_cuda": (t_r / t_o) if t_r else None}
    if note:
        ent["note"] = note
    ent.update(extra)
    return ent


def refcuda():
    """tests/_refcuda (loader of oracle/_ref/nerfacc_cuda.so) or None when it was not built."""
    try:
        from tests import _refcuda as R
    except Exception:
        return None
    return R if R.available() else None


def refpkg():
    """The reference's own Python package on its own kernels (tests/_refpkg) or None."""
    try:
        from tests import _refpkg as P
    except Exception:
        return None
    return P.load() if P.available() else None


# ------------------------------------------------------------------ reference compositions
class _RefExclusiveSum(torch.autograd.Function):
    """The reference's _ExclusiveSum (nerfacc/scan.py:218-242) on its own kernels."""

    @staticmethod
    def forward(ctx, R, x, pi):
        ctx.R, ctx.pi = R, pi
        return R.scan(x, pi, "sum", True)

    @staticmethod
    def backward(ctx, g):
        return None, ctx.R.scan(g.contiguous(), ctx.pi, "sum", True, backward=True), None


class _RefExclusiveProd(torch.autograd.Function):
    """The reference's _ExclusiveProd (nerfacc/scan.py:267-288) on its own kernels."""

    @staticmethod
    def forward(ctx, R, x, pi):
        y = R.scan(x, pi, "prod", True)
        ctx.R, ctx.pi = R, pi
        ctx.save_for_backward(x, y)
        return y

    @staticmethod
    def backward(ctx, g):
        x, y = ctx.saved_tensors
        return None, ctx.R.prod_backward(g.contiguous(), x, y, ctx.pi, True), None


def ref_weight_from_density(R, ts, te, sig, pi):
    """nerfacc/volrend.py:209-264 + 312-362 composed as the reference does."""
    sd = sig * (te - ts)
    alphas = 1.0 - torch.exp(-sd)
    trans = torch.exp(-_RefExclusiveSum.apply(R, sd, pi))
    return trans * alphas, trans, alphas


def ref_weight_from_alpha(R, alphas, pi):
    """nerfacc/volrend.py:161-206 + 267-309."""
    trans = _RefExclusiveProd.apply(R, 1.0 - alphas, pi)
    return trans * alphas, trans


def ref_accumulate(w, v, ri, n_rays):
    """nerfacc/volrend.py:532-547."""
    src = w[:, None] * v if v is not None else w[:, None]
    out = torch.zeros((n_rays, src.shape[-1]), device=w.device, dtype=src.dtype)
    out.index_add_(0, ri, src)
    return out


def ref_pack_info(ri, n_rays):
    """nerfacc/pack.py:38-46, run at every render_* call of the reference."""
    cnts = torch.zeros((n_rays,), device=ri.device, dtype=torch.long)
    cnts.index_add_(0, ri, torch.ones_like(ri))
    return torch.stack([cnts.cumsum(0) - cnts, cnts], -1)


def ref_sampling(R, o, d, binaries, aabbs, near, far, step, cone=0.0):
    """OccGridEstimator.sampling without callbacks on the reference kernels (occ_grid.py:163-190)."""
    n = o.shape[0]
    near_t = torch.full((n,), float(near), device=o.device)
    far_t = torch.full((n,), float(far), device=o.device)
    iv, sm, _ = R.traverse_grids(o, d, binaries, aabbs, near_t, far_t, step, cone)
    return sm["ray_indices"], iv["vals"][iv["is_left"]], iv["vals"][iv["is_right"]], sm["packed_info"]


# ------------------------------------------------------------------ error metric
def err_vs(x: torch.Tensor, truth: torch.Tensor) -> float:
    """max_i |x_i - t_i| / (|t_i| + 1e-2 * rms(t)): a relative error that measures elements far
    below the array's scale against that scale (1 % of the RMS) instead of against themselves."""
    t = truth.double()
    scale = float(t.pow(2).mean().sqrt().clamp_min(1e-30))
    return float(((x.double() - t).abs() / (t.abs() + 1e-2 * scale)).max())


def rms_err(x: torch.Tensor, truth: torch.Tensor) -> float:
    """rms(x - t) / rms(t): the error of the array as a whole (a max over 10^7 elements is an
    extreme-value statistic and moves by tens of percent between two equally accurate methods)."""
    t = truth.double()
    return float(((x.double() - t).pow(2).mean().sqrt() / t.pow(2).mean().sqrt().clamp_min(1e-300)))


def as_close_as_reference(e_ours: float, e_ref: float, rms_ours: float, rms_ref: float, tol: float = 1e-5) -> bool:
    """The float acceptance rule: within `tol` (north_star: 1e-5 relative) of float64, or as
    close to float64 as the reference's own kernels are — max error within 2x of theirs and RMS
    error within 1.25x of theirs (fp32 formulas such as 1 - exp(-x) carry errors far above 1e-5
    for small x whichever kernel evaluates them)."""
    return (e_ours <= max(tol, 2.0 * e_ref)) and (rms_ours <= max(0.1 * tol, 1.25 * rms_ref))


def _f64_step(ts, te, ri, pi, sig, rgb, g_c, g_o, g_d, n_rays):
    """The headline step's arithmetic in float64 (torch, global cumsum minus the ray's offset)."""
    ts, te, sig, rgb = ts.double(), te.double(), sig.double().requires_grad_(True), rgb.double().requires_grad_(True)
    sd = sig * (te - ts)
    cs = torch.cumsum(sd, 0)
    excl = cs - sd
    first = pi[:, 0].clamp_max(max(sd.numel() - 1, 0))
    base = excl[first]  # value of the global exclusive sum at each ray's first sample
    trans = torch.exp(-(excl - base[ri]))
    alphas = 1.0 - torch.exp(-sd)
    w = trans * alphas
    col = torch.zeros((n_rays, 3), dtype=torch.float64, device=ts.device).index_add_(0, ri, w[:, None] * rgb)
    opa = torch.zeros((n_rays, 1), dtype=torch.float64, device=ts.device).index_add_(0, ri, w[:, None])
    dep = torch.zeros((n_rays, 1), dtype=torch.float64, device=ts.device).index_add_(
        0, ri, (w * (ts + te) * 0.5)[:, None])
    dep = dep / opa.clamp_min(torch.finfo(torch.float32).eps)
    gs, gr = torch.autograd.grad([col, opa, dep], [sig, rgb], [g_c.double(), g_o.double(), g_d.double()])
    return dict(weights=w.detach(), colors=col.detach(), opacities=opa.detach(), depths=dep.detach(),
                g_sigmas=gs, g_rgbs=gr)


def preflight_parity(nf, est, o, d, step_size, sigmas, rgbs, g_c, g_o, g_d, tol=1e-5):
    """The headline step at its full size against the reference's own CUDA kernels, outside the
    timed region: sampling outputs bit-exact; every float output and gradient measured against a
    float64 evaluation, ours next to the reference composition (same metric), and required to be
    within `tol` or as close to float64 as the reference is."""
    R = refcuda()
    if R is None:
        return {"status": "skipped", "why": "oracle/_ref/nerfacc_cuda.so not present"}
    n_rays = o.shape[0]
    out = {"status": "ok", "rays": n_rays, "checker": "oracle/_ref/nerfacc_cuda.so (unmodified reference kernels)"}
    ri, ts, te = est.sampling(o, d, render_step_size=step_size)
    pi = nf.pack_info(ri, n_rays)
    r_ri, r_ts, r_te, r_pi = ref_sampling(R, o, d, est.binaries, est.aabbs, 0.0, 1e10, step_size)
    exact = {"ray_indices": torch.equal(ri, r_ri), "t_starts": torch.equal(ts, r_ts),
             "t_ends": torch.equal(te, r_te), "packed_info": torch.equal(pi, r_pi)}
    out["samples"] = int(ri.numel())
    out["sampling"] = "bit-exact" if all(exact.values()) else {k: v for k, v in exact.items() if not v}
    del r_ts, r_te
    sg = sigmas.detach().clone().requires_grad_(True)
    rg = rgbs.detach().clone().requires_grad_(True)
    c, op, dp, ex = nf.rendering(ts, te, ri, n_rays, rgb_sigma_fn=lambda a, b, r: (rg, sg))
    gs, gr = torch.autograd.grad([c, op, dp], [sg, rg], [g_c, g_o, g_d])
    ours = dict(weights=ex["weights"].detach(), colors=c.detach(), opacities=op.detach(), depths=dp.detach(),
                g_sigmas=gs, g_rgbs=gr)
    del c, op, dp, ex
    sg2 = sigmas.detach().clone().requires_grad_(True)
    rg2 = rgbs.detach().clone().requires_grad_(True)
    wts, _, _ = ref_weight_from_density(R, ts, te, sg2, ref_pack_info(r_ri, n_rays))
    rc = ref_accumulate(wts, rg2, r_ri, n_rays)
    ro = ref_accumulate(wts, None, r_ri, n_rays)
    rd = ref_accumulate(wts, ((ts + te) * 0.5)[:, None], r_ri, n_rays)
    rd = rd / ro.clamp_min(torch.finfo(torch.float32).eps)
    rgs, rgr = torch.autograd.grad([rc, ro, rd], [sg2, rg2], [g_c, g_o, g_d])
    ref = dict(weights=wts.detach(), colors=rc.detach(), opacities=ro.detach(), depths=rd.detach(),
               g_sigmas=rgs, g_rgbs=rgr)
    del wts, rc, ro, rd
    truth = _f64_step(ts, te, ri, pi, sigmas.detach(), rgbs.detach(), g_c, g_o, g_d, n_rays)
    errs, ok = {}, all(exact.values())
    for k in ours:
        e_o, e_r = err_vs(ours[k], truth[k]), err_vs(ref[k], truth[k])
        r_o, r_r = rms_err(ours[k], truth[k]), rms_err(ref[k], truth[k])
        errs[k] = {"ours_vs_f64": e_o, "ref_cuda_vs_f64": e_r, "ours_vs_ref_cuda": err_vs(ours[k], ref[k]),
                   "rms_ours_vs_f64": r_o, "rms_ref_cuda_vs_f64": r_r}
        good = as_close_as_reference(e_o, e_r, r_o, r_r, tol)
        if not good:
            errs[k]["FAILED"] = True
        ok &= good
    out["float_errors"] = errs
    out["metric"] = ("max |x - t| / (|t| + 1e-2 rms(t)) and rms(x - t) / rms(t) against float64; pass: max <= "
                     "max(1e-5, 2 x the reference kernels') and rms <= max(1e-6, 1.25 x the reference kernels')")
    if not ok:
        out["status"] = "FAILED"
    return out


# ------------------------------------------------------------------ cfg3
def cfg3_counts(n_rays, total, seed=1234):
    """c_r = clamp(round(exp(N(ln 12, 0.9^2))), 0, 2048), 20 % of rays empty,
    rescaled so the sum is `total` within 0.1 %."""
    rng = np.random.default_rng(seed)
    base = np.exp(rng.normal(np.log(12.0), 0.9, n_rays))
    empty = rng.random(n_rays) < 0.2
    scale = 1.0
    for _ in range(30):
        c = np.clip(np.round(base * scale), 0, 2048).astype(np.int64)
        c[empty] = 0
        s = int(c.sum())
        if abs(s - total) <= total * 1e-3:
            break
        scale *= total / max(s, 1)
    return c


def bench_cfg3(nf, dev, peak, R, n_rays=1 << 20, total=1 << 24):
    from nerfacc_b200._packing import mark_canonical

    c = cfg3_counts(n_rays, total)
    cnt = torch.from_numpy(c).to(dev)
    pi = torch.stack([torch.cumsum(cnt, 0) - cnt, cnt], -1).contiguous()
    S = int(cnt.sum())
    ri = torch.repeat_interleave(torch.arange(n_rays, device=dev), cnt)
    pi = nf.pack_info(ri, n_rays)  # the product's own (canonical, dense) packed_info
    g0 = torch.Generator(device=dev).manual_seed(1234)
    alphas = torch.rand(S, device=dev, generator=g0) ** 4
    g = torch.rand(S, device=dev, generator=g0)
    ts = torch.rand(S, device=dev, generator=g0)
    te = ts + 0.01
    res = {"samples": S, "rays": n_rays, "max_count": int(c.max())}
    a1 = alphas.clone().requires_grad_(True)

    def ours_fb():
        w, _ = nf.render_weight_from_alpha(a1, packed_info=pi)
        return torch.autograd.grad(w, a1, g)[0]

    def ref_fb():
        w, _ = ref_weight_from_alpha(R, a1, pi)
        return torch.autograd.grad(w, a1, g)[0]

    res["weights_fwd_bwd"] = _entry(28 * S + 32 * n_rays, ours_fb, ref_fb if R else None, peak,
                                    note="render_weight_from_alpha fwd + (w*g).sum().backward(); 12S+16R fwd, 16S+16R bwd")
    if R:
        res["weights_fwd_bwd"]["err_grad_ours_vs_ref_cuda"] = err_vs(ours_fb(), ref_fb())

    # pack_info from ray_indices (a fresh view every call: the identity-keyed recall of a known
    # packed_info must miss; the view itself moves no data)
    P = refpkg()
    res["pack_info"] = _entry(8 * S + 16 * n_rays, lambda: nf.pack_info(ri[:], n_rays),
                              (lambda: P.pack_info(ri, n_rays)) if P is not None else None, peak,
                              note="pack_info(ray_indices): read 8 B/sample, write (start, count) per ray; the reference "
                                   "(pack.py:43-46) is index_add_ + cumsum + stack in torch")
    if P is not None:
        res["pack_info"]["equal_to_ref"] = bool(torch.equal(nf.pack_info(ri[:], n_rays), P.pack_info(ri, n_rays).to(torch.long)))

    eps, thre = 1e-4, 1e-2

    from nerfacc_b200 import _occ_ops

    def ours_vis(want_mask=False):
        with torch.no_grad():
            return _occ_ops.visibility_compact(False, alphas, ts, te, pi, ri, eps, thre, want_mask=want_mask)

    def ref_vis():
        with torch.no_grad():
            trans = R.scan(1.0 - alphas, pi, "prod", True)
            vis = (trans >= eps) & (alphas >= thre)
            return ri[vis], ts[vis], te[vis]

    k = int(ours_vis()[0].numel())
    res["kept"] = k
    res["visibility_compaction"] = _entry(
        4 * S + 16 * S + 16 * k + 32 * n_rays, ours_vis, ref_vis if R else None, peak,
        note="render_visibility_from_alpha(early_stop_eps=1e-4, alpha_thre=1e-2) + compaction of "
             "(ray_indices, t_starts, t_ends), occ_grid.py:199-220")
    if R:
        a, b = ours_vis()[:3], ref_vis()
        res["visibility_compaction"]["equal_to_ref_cuda"] = all(torch.equal(x, y) for x, y in zip(a, b))

    # the whole cfg3 sequence: visibility -> compaction -> weights fwd+bwd on the survivors
    def ours_seq():
        r2, t0, t1, pi2, m = ours_vis(want_mask=True)
        a2 = alphas[m].requires_grad_(True)
        mark_canonical(pi2, total=r2.numel())
        w, _ = nf.render_weight_from_alpha(a2, packed_info=pi2)
        return torch.autograd.grad(w, a2, g[: w.numel()])[0]

    def ref_seq():
        with torch.no_grad():
            trans = R.scan(1.0 - alphas, pi, "prod", True)
            vis = (trans >= eps) & (alphas >= thre)
            r2, t0, t1 = ri[vis], ts[vis], te[vis]
        a2 = alphas[vis].requires_grad_(True)
        w, _ = ref_weight_from_alpha(R, a2, ref_pack_info(r2, n_rays))
        return torch.autograd.grad(w, a2, g[: w.numel()])[0]

    try:
        res["visibility_compaction_weights"] = _entry(
            (4 + 16) * S + 16 * k + 32 * n_rays + 4 * S + 4 * k + 28 * k + 32 * n_rays, ours_seq,
            ref_seq if R else None, peak,
            note="visibility -> compaction -> alphas[vis] -> render_weight_from_alpha fwd+bwd on the survivors")
    except Exception as e:  # pragma: no cover - reported, not fatal
        res["visibility_compaction_weights"] = {"error": repr(e)}
    return res


# ------------------------------------------------------------------ cfg4
CFG4 = dict(roi=[-1.0, -1.0, -1.0, 1.0, 1.0, 1.0], levels=4, res=128, near=0.2, far=1e10, step=1e-3,
            cone=0.004, shell=(0.8, 0.9))


def cfg4_binaries(dev, seed=1234):
    """binaries[l] = shell (cell-centre radius in [0.8, 0.9] x 2^l, i.e. the same shell in level-l
    coordinates) union 1 % random cells."""
    res, m = CFG4["res"], CFG4["levels"]
    lin = ((torch.arange(res, device=dev, dtype=torch.float64) + 0.5) / res * 2.0 - 1.0)
    X, Y, Z = torch.meshgrid(lin, lin, lin, indexing="ij")
    rad = torch.sqrt(X * X + Y * Y + Z * Z)  # in units of the level's half extent
    shell = (rad >= CFG4["shell"][0]) & (rad <= CFG4["shell"][1])
    g = torch.Generator(device=dev).manual_seed(seed)
    return torch.stack([shell | (torch.rand((res, res, res), device=dev, generator=g) < 0.01) for _ in range(m)])


def cfg4_rays(n_rays, seed, dev):
    rng = np.random.default_rng(seed)
    o = rng.standard_normal((n_rays, 3))
    o = 4.0 * o / np.linalg.norm(o, axis=-1, keepdims=True)
    tgt = rng.standard_normal((n_rays, 3))
    tgt = tgt / np.linalg.norm(tgt, axis=-1, keepdims=True) * 1.0 * rng.random((n_rays, 1)) ** (1 / 3)
    d = tgt - o
    d /= np.linalg.norm(d, axis=-1, keepdims=True)
    return torch.from_numpy(o.astype(np.float32)).to(dev), torch.from_numpy(d.astype(np.float32)).to(dev)


def cfg4_occ_eval_fn(x):
    """occ_eval_fn of SURVEY §8d cfg4, made smooth in x so that jittered points inside one cell
    give different values: step * 50 * bump(|x| around the level-0 shell) + a slow radial tail."""
    r = x.norm(dim=-1, keepdim=True)
    return CFG4["step"] * 50.0 * (torch.exp(-((r - 0.85) / 0.05) ** 2) + 0.02 / (1.0 + r))


def bench_cfg4(nf, dev, peak, R, n_rays=1 << 20):
    est = nf.OccGridEstimator(CFG4["roi"], resolution=CFG4["res"], levels=CFG4["levels"]).to(dev)
    est.binaries = cfg4_binaries(dev)
    o, d = cfg4_rays(n_rays, 4321, dev)
    kw = dict(render_step_size=CFG4["step"], cone_angle=CFG4["cone"], near_plane=CFG4["near"], far_plane=CFG4["far"])
    ri, ts, te = est.sampling(o, d, **kw)
    S = int(ri.numel())
    res = {"rays": n_rays, "samples": S, "levels": CFG4["levels"]}
    B = est.binaries.numel() // 8
    ent = _entry(16 * S + 48 * n_rays + B, lambda: est.sampling(o, d, **kw),
                 (lambda: ref_sampling(R, o, d, est.binaries, est.aabbs, CFG4["near"], CFG4["far"], CFG4["step"],
                                       CFG4["cone"])) if R else None, peak,
                 note="OccGridEstimator.sampling, 4 nested 128^3 levels, cone stepping (dt = max(1e-3, 0.004 t))")
    if R:
        r_ri, r_ts, r_te, _ = ref_sampling(R, o, d, est.binaries, est.aabbs, CFG4["near"], CFG4["far"], CFG4["step"],
                                           CFG4["cone"])
        ent["equal_to_ref_cuda"] = bool(torch.equal(ri, r_ri) and torch.equal(ts, r_ts) and torch.equal(te, r_te))
        del r_ri, r_ts, r_te
    res["sampling"] = ent
    del ri, ts, te
    # update_every_n_steps: warm-up (all 4 * 2^21 cells) and steady state (<= 2 * res^3/4 per level)
    est.train()

    def upd(step):
        est._update(step, cfg4_occ_eval_fn, occ_thre=1e-2, ema_decay=0.95, warmup_steps=256)

    cells = est.levels * est.cells_per_lvl
    t_w = timeit(lambda: upd(0), warm=2, groups=3, reps=3)
    res["update_warmup"] = {"ms": t_w, "cells": cells, "bytes": 32 * cells + 5 * cells,
                            "gbs": 37 * cells / (t_w * 1e-3) / 1e9, "frac": 37 * cells / (t_w * 1e-3) / 1e9 / peak,
                            "note": "per cell: 8 idx + 12 points + 4 occ + 8 occs RMW, then binarise 4 read + 1 write"}
    t_s = timeit(lambda: upd(512), warm=2, groups=3, reps=3)
    res["update_steady"] = {"ms": t_s, "cells_sampled_at_most": 2 * (est.cells_per_lvl // 4) * est.levels}
    return res


def cfg4_sharded_update_check(nf, dev, dist, rank, world):
    """BASELINE cfg4's collective, on NCCL: OccGridEstimator._update through shard_updates()
    at one warm-up step and two steady-state steps on every rank, against the same updates done
    un-sharded by rank 0 (same private seeds): `occs` and `binaries` must be torch.equal on all
    ranks and to the un-sharded copy.  Returns the `update` object of the JSON line."""
    est = nf.OccGridEstimator(CFG4["roi"], resolution=CFG4["res"], levels=CFG4["levels"]).to(dev)
    est.train()
    est.shard_updates(None, seed=11)
    single = None
    if rank == 0:
        single = nf.OccGridEstimator(CFG4["roi"], resolution=CFG4["res"], levels=CFG4["levels"]).to(dev)
        single.train()
        single.shard_updates(None, seed=11, _world_rank=(1, 0))
    steps = (0, 256, 272)
    ok, ms = True, {}
    for step in steps:
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        est._update(step, cfg4_occ_eval_fn, occ_thre=1e-2, ema_decay=0.95, warmup_steps=256)
        e1.record()
        torch.cuda.synchronize()
        ms[str(step)] = e0.elapsed_time(e1)
        if single is not None:
            single._update(step, cfg4_occ_eval_fn, occ_thre=1e-2, ema_decay=0.95, warmup_steps=256)
        # every rank against rank 0's sharded grid, rank 0 against its un-sharded copy
        ref_occs = est.occs.clone()
        ref_bins = est.binaries.clone()
        dist.broadcast(ref_occs, 0)
        bb = ref_bins.to(torch.uint8)
        dist.broadcast(bb, 0)
        same = torch.equal(est.occs, ref_occs) and torch.equal(est.binaries.to(torch.uint8), bb)
        if single is not None:
            same = same and torch.equal(est.occs, single.occs) and torch.equal(est.binaries, single.binaries)
        flag = torch.tensor([1 if same else 0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        ok &= bool(flag.item())
    # the collective alone: MAX all-reduce of the 32 MiB fp32 `occs`
    buf = est.occs.clone()
    for _ in range(3):
        dist.all_reduce(buf, op=dist.ReduceOp.MAX)
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        dist.all_reduce(buf, op=dist.ReduceOp.MAX)
    e1.record()
    torch.cuda.synchronize()
    ar = torch.tensor([e0.elapsed_time(e1) / 10], device=dev, dtype=torch.float64)
    dist.all_reduce(ar, op=dist.ReduceOp.MAX)
    t = torch.tensor([ms[str(s)] for s in steps], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    occupied = int(est.binaries.sum().item())
    return {"parity": "bit-exact" if ok else "MISMATCH", "config": "cfg4: 4 x 128^3 grid, roi [-1,1]^3",
            "steps_checked": list(steps), "against": "all ranks equal + rank 0's un-sharded update from the same seeds",
            "ms": {str(s): float(v) for s, v in zip(steps, t.tolist())}, "allreduce_ms": float(ar.item()),
            "bytes": int(est.occs.numel() * 4), "occupied_cells": occupied, "world": world}


# ------------------------------------------------------------------ cfg5
def cfg5_sigma_fn(n_rays, dev, seed=1234):
    """Analytic proposal density: two Gaussian bumps in t with per-ray random centres."""
    g = torch.Generator(device=dev).manual_seed(seed)
    c1 = 1.0 + 3.0 * torch.rand((n_rays, 1), device=dev, generator=g)
    c2 = 5.0 + 20.0 * torch.rand((n_rays, 1), device=dev, generator=g)
    scale = torch.nn.Parameter(torch.tensor(1.0, device=dev))

    def fn(t_starts, t_ends):
        t = (t_starts + t_ends) * 0.5
        return scale * (8.0 * torch.exp(-((t - c1) / 0.3) ** 2) + 0.5 * torch.exp(-((t - c2) / 3.0) ** 2))

    return fn, scale


def bench_cfg5(nf, dev, peak, R, n_rays=1 << 20):
    from nerfacc_b200.estimators.prop_net import _pdf_loss, _transform_stot, _cdf_from_trans

    res = {"rays": n_rays, "levels": "2 -> 257 -> 97 -> 49 edges"}
    near, far = 0.2, 1e3
    fn, scale = cfg5_sigma_fn(n_rays, dev)
    cd = torch.tensor([0.0, 1.0], device=dev).repeat(n_rays, 1)
    lv = cd.clone()
    kept = []
    for n_out in (256, 96, 48):
        n_in = lv.shape[1]
        name = f"importance_sampling_{n_in}_to_{n_out}"
        res[name] = _entry(
            8 * n_in * n_rays + (4 * n_out + 4) * n_rays,
            lambda lv=lv, cd=cd, n_out=n_out: nf.importance_sampling(nf.RayIntervals(vals=lv), cd, n_out, _samples=False),
            (lambda lv=lv, cd=cd, n_out=n_out: R.importance_sampling(lv, cd, n_out)) if R else None, peak,
            note="intervals only, as PropNetEstimator.sampling calls it (the reference writes the centres too)")
        iv, _ = nf.importance_sampling(nf.RayIntervals(vals=lv), cd, n_out, _samples=False)
        if R:
            r_iv, _ = R.importance_sampling(lv, cd, n_out)
            res[name]["equal_to_ref_cuda"] = bool(torch.equal(iv.vals, r_iv))
            del r_iv
        lv = iv.vals
        if n_out != 48:
            with torch.no_grad():
                t = _transform_stot("lindisp", lv, near, far)
                trans, _ = nf.render_transmittance_from_density(t[..., :-1], t[..., 1:], fn(t[..., :-1], t[..., 1:]))
                cd = _cdf_from_trans(trans)
            kept.append((lv, cd))
    q = lv  # [R, 49]
    key, key_cdf = kept[0]
    res["searchsorted_49_in_257"] = _entry(
        4 * (q.shape[1] + key.shape[1]) * n_rays + 16 * q.shape[1] * n_rays,
        lambda: nf.searchsorted(nf.RayIntervals(vals=key), nf.RayIntervals(vals=q)),
        (lambda: R.searchsorted(key, q)) if R else None, peak)
    if R:
        a, b = nf.searchsorted(nf.RayIntervals(vals=key), nf.RayIntervals(vals=q)), R.searchsorted(key, q)
        res["searchsorted_49_in_257"]["equal_to_ref_cuda"] = bool(torch.equal(a[0], b[0]) and torch.equal(a[1], b[1]))
    # _pdf_loss forward + backward into the key level's cdf (scatter-add), prop_net.py:232-256
    qc = torch.rand((n_rays, 49), device=dev).cumsum(-1)
    qc = qc / qc[:, -1:]
    kc = key_cdf.clone().requires_grad_(True)

    def loss_fb():
        loss = _pdf_loss(nf.RayIntervals(vals=q), qc, nf.RayIntervals(vals=key), kc).mean()
        return torch.autograd.grad(loss, kc)[0]

    t = timeit(loss_fb)
    res["pdf_loss_fwd_bwd_257"] = {"ms": t, "note": "_pdf_loss(49 query edges, 257 key edges).mean() + backward to cdfs_key"}
    del kc, qc, kept, key, key_cdf, q, lv, cd
    torch.cuda.empty_cache()
    # ---- the estimator end to end, next to the reference's own PropNetEstimator (baseline/_ref)
    ref = refpkg()
    args = ([256, 96], 48, n_rays, near, far, "lindisp", False)
    est = nf.PropNetEstimator().to(dev)
    r_est = ref.PropNetEstimator().to(dev) if ref is not None else None
    # (a) analytic proposal densities (two Gaussian bumps per ray: ~15 elementwise torch kernels per level,
    # the user's side of the callback, identical for both estimators)
    t_o = timeit(lambda: est.sampling([fn, fn], *args, False), warm=2, groups=3, reps=3)
    t_r = timeit(lambda: r_est.sampling([fn, fn], *args, False), warm=2, groups=3, reps=3) if r_est else None
    res["propnet_sampling"] = {"ms": t_o, "ref_pkg_ms": t_r, "speedup_vs_ref_pkg": (t_r / t_o) if t_r else None,
                               "note": "PropNetEstimator.sampling(prop_samples=[256, 96], num_samples=48), lindisp, "
                                       "analytic proposal densities, requires_grad=False; ref_pkg = the reference's "
                                       "PropNetEstimator on the reference kernels with the same callbacks"}
    # (b) the estimator's own work: proposal networks that cost nothing (they return stored tensors)
    stored = {256: torch.rand((n_rays, 256), device=dev), 96: torch.rand((n_rays, 96), device=dev)}
    free = lambda a, b: stored[a.shape[-1]]
    t_o = timeit(lambda: est.sampling([free, free], *args, False), warm=2, groups=3, reps=3)
    t_k, names = native_ms(lambda: est.sampling([free, free], *args, False), reps=3)
    t_r = timeit(lambda: r_est.sampling([free, free], *args, False), warm=2, groups=3, reps=3) if r_est else None
    by = n_rays * 4 * ((2 * 2 + 2 * 257) + (2 * 257 + 256 + 257) + (2 * 257 + 2 * 97) + (2 * 97 + 96 + 97) + (2 * 97 + 2 * 49))
    res["propnet_sampling_own_work"] = {
        "ms": t_o, "kernel_ms": t_k, "native_calls": names, "bytes": by, "gbs": by / (t_k * 1e-3) / 1e9,
        "frac": by / (t_k * 1e-3) / 1e9 / peak, "ref_pkg_ms": t_r, "speedup_vs_ref_pkg": (t_r / t_o) if t_r else None,
        "note": "the same call with proposal networks that return stored tensors: what the estimator itself costs; "
                "bytes = three resampling levels (read edges+cdf, write s and t edges) + two cdf constructions "
                "(read t edges + sigmas, write cdf)"}
    # (c) training side: sampling with requires_grad, then compute_loss and backward to the proposal densities
    n_tr = n_rays // 4
    sg = {256: torch.rand((n_tr, 256), device=dev, requires_grad=True),
          96: torch.rand((n_tr, 96), device=dev, requires_grad=True)}
    learn = lambda a, b: sg[a.shape[-1]]
    trans = torch.rand((n_tr, 48), device=dev).cumsum(-1).neg().exp()
    targs = ([256, 96], 48, n_tr, near, far, "lindisp", False)

    def train(e):
        e.sampling([learn, learn], *targs, True)
        loss = e.compute_loss(trans)
        return torch.autograd.grad(loss, list(sg.values()))

    t_o = timeit(lambda: train(est), warm=2, groups=3, reps=3)
    t_k, names = native_ms(lambda: train(est), reps=3)
    t_r = timeit(lambda: train(r_est), warm=2, groups=3, reps=3) if r_est else None
    res["propnet_train_step"] = {
        "rays": n_tr, "ms": t_o, "kernel_ms": t_k, "native_calls": names, "ref_pkg_ms": t_r,
        "speedup_vs_ref_pkg": (t_r / t_o) if t_r else None,
        "note": "sampling(requires_grad=True) + compute_loss + backward to the two levels' densities (stored tensors)"}
    return res


# ------------------------------------------------------------------ test-time loop
def spelled_out_loop(pkg, est, o, d, fn, step, cone, early_stop_eps, alpha_thre, max_samples=1024):
    """examples/utils.py:252-425 (render_image_with_occgrid_test) call for call, on package `pkg`
    (this library, or the reference package on the reference kernels)."""
    n = o.shape[0]
    dev = o.device
    acc_ = getattr(pkg, "accumulate_along_rays_", None) or pkg.volrend.accumulate_along_rays_
    opacity = torch.zeros((n, 1), device=dev)
    depth = torch.zeros((n, 1), device=dev)
    rgb = torch.zeros((n, 3), device=dev)
    ray_mask = torch.ones(n, device=dev).bool()
    min_samples = 1 if cone == 0 else 4
    iter_samples = total_samples = 0
    near_planes = torch.full_like(o[..., 0], fill_value=0.0)
    far_planes = torch.full_like(o[..., 0], fill_value=1e10)
    t_mins, t_maxs, hits = pkg.ray_aabb_intersect(o, d, est.aabbs)
    m = est.binaries.size(0)
    if m > 1:
        t_sorted, t_indices = torch.sort(torch.cat([t_mins, t_maxs], -1), -1)
    else:
        t_sorted = torch.cat([t_mins, t_maxs], -1)
        t_indices = torch.arange(0, m * 2, device=dev, dtype=torch.int64).expand(n, m * 2)
    opc_thre = 1 - early_stop_eps
    while iter_samples < max_samples:
        n_alive = ray_mask.sum().item()
        if n_alive == 0:
            break
        n_samples = max(min(n // n_alive, 64), min_samples)
        iter_samples += n_samples
        intervals, samples, termination_planes = pkg.traverse_grids(
            o, d, est.binaries, est.aabbs, near_planes, far_planes, step, cone, n_samples, True, ray_mask,
            t_sorted, t_indices, hits)
        t_starts = intervals.vals[intervals.is_left]
        t_ends = intervals.vals[intervals.is_right]
        ray_indices = samples.ray_indices[samples.is_valid]
        packed_info = samples.packed_info
        rgbs, sigmas = fn(t_starts, t_ends, ray_indices)
        weights, _, alphas = pkg.render_weight_from_density(
            t_starts, t_ends, sigmas, ray_indices=ray_indices, n_rays=n,
            prefix_trans=1 - opacity[ray_indices].squeeze(-1))
        if alpha_thre > 0:
            vis = alphas >= alpha_thre
            ray_indices, rgbs, weights, t_starts, t_ends = (ray_indices[vis], rgbs[vis], weights[vis],
                                                            t_starts[vis], t_ends[vis])
        acc_(weights, values=rgbs, ray_indices=ray_indices, outputs=rgb)
        acc_(weights, values=None, ray_indices=ray_indices, outputs=opacity)
        acc_(weights, values=(t_starts + t_ends)[..., None] / 2.0, ray_indices=ray_indices,
                                   outputs=depth)
        near_planes = termination_planes
        ray_mask = torch.logical_and(opacity.view(-1) <= opc_thre, packed_info[:, 1] == n_samples)
        total_samples += ray_indices.shape[0]
    depth = depth / opacity.clamp_min(torch.finfo(torch.float32).eps)
    return rgb, opacity, depth, total_samples


def bench_chunked(nf, dev, peak, n_rays=1 << 18):
    """The test-time loop on the cfg2 scene (thin shell, step 5e-3) with an analytic field
    (sigma = 50 inside the shell): this library's fused loop, the same loop spelled with this
    library's public API, and the reference package on the reference kernels."""
    import bench as B

    est = nf.OccGridEstimator(B.AABB.tolist(), resolution=B.RES, levels=1).to(dev)
    est.binaries = torch.from_numpy(B.make_grid()).to(dev)
    o_np, d_np = B.make_rays(n_rays, 77)
    o, d = torch.from_numpy(o_np).to(dev), torch.from_numpy(d_np).to(dev)

    def fn(t_starts, t_ends, ray_indices):
        pos = nf.sample_positions(o, d, ray_indices, t_starts, t_ends)
        rad = pos.norm(dim=-1)
        return torch.sigmoid(pos * 3.0), 50.0 * ((rad >= 0.8) & (rad <= 0.9)).float()

    kw = dict(step=B.STEP_SIZE, cone=0.0, early_stop_eps=1e-4, alpha_thre=0.0)
    fused = lambda: nf.render_image_with_occgrid_test(1024, fn, est, o, d, render_step_size=B.STEP_SIZE,
                                                      early_stop_eps=1e-4)
    rgb, opac, depth, total = fused()
    iters = []
    nf.render_image_with_occgrid_test(1024, fn, est, o, d, render_step_size=B.STEP_SIZE, early_stop_eps=1e-4,
                                      _on_chunk=lambda c: iters.append(c["k"]))
    res = {"rays": n_rays, "samples_marched": int(total), "iterations": len(iters)}
    t_f = timeit(fused, warm=1, groups=3, reps=1)
    t_k, names = native_ms(fused, reps=1)
    t_s = timeit(lambda: spelled_out_loop(nf, est, o, d, fn, **kw), warm=1, groups=3, reps=1)
    res["fused_loop"] = {"ms": t_f, "kernel_ms": t_k, "native_calls": names, "host_syncs_per_iteration": 1}
    res["public_api_loop"] = {"ms": t_s, "note": "examples/utils.py:252-425 spelled with this library's public API"}
    ref = refpkg()
    if ref is not None:
        r_est = ref.OccGridEstimator(B.AABB.tolist(), resolution=B.RES, levels=1).to(dev)
        r_est.binaries = est.binaries.clone()
        t_r = timeit(lambda: spelled_out_loop(ref, r_est, o, d, fn, **kw), warm=1, groups=3, reps=1)
        r2 = spelled_out_loop(ref, r_est, o, d, fn, **kw)
        same = (opac - r2[1]).abs().view(-1) <= 1e-4
        res["reference_loop"] = {"ms": t_r, "samples_marched": int(r2[3]),
                                 "note": "the same loop on the reference package and kernels (baseline/_ref, oracle/_ref)"}
        res["fused_loop"]["speedup_vs_reference_loop"] = t_r / t_f
        res["fused_loop"]["rays_with_equal_opacity"] = float(same.float().mean())
        res["fused_loop"]["max_rgb_diff_on_those"] = float((rgb - r2[0]).abs()[same].max())
    return res


# ------------------------------------------------------------------ cfg1 (CPU)
def cpu_cfg1(steps=8, warmup=2):
    """BASELINE.json configs[0] on the host cores: the reference's batched PyTorch path —
    render_weight_from_density (nerfacc/volrend.py:312-362 -> render_transmittance_from_density
    :209-264 -> batched exclusive_sum = torch.cumsum of the shifted input, nerfacc/scan.py:84-91)
    and accumulate_along_rays (:532-547, the batched `sum(dim=-2)` branch), forward + backward,
    restated in torch (the reference package does not travel to the GPU box)."""
    g = torch.Generator().manual_seed(1234)
    n_rays, n = 4096, 128
    t = torch.linspace(0.0, 1.0, n + 1).expand(n_rays, n + 1)
    ts, te = t[:, :-1], t[:, 1:]
    sigmas = (10.0 * torch.rand((n_rays, n), generator=g)).requires_grad_(True)
    rgbs = torch.rand((n_rays, n, 3), generator=g).requires_grad_(True)

    def step():
        sigmas.grad = None
        rgbs.grad = None
        sd = sigmas * (te - ts)
        alphas = 1.0 - torch.exp(-sd)
        excl = torch.cumsum(torch.cat([torch.zeros_like(sd[..., :1]), sd[..., :-1]], dim=-1), dim=-1)
        trans = torch.exp(-excl)
        weights = trans * alphas
        colors = (weights[..., None] * rgbs).sum(dim=-2)
        colors.sum().backward()

    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    return {"value": n_rays * n / dt / 1e9, "unit": "Gsamples/s", "ms_per_step": dt * 1e3,
            "threads": torch.get_num_threads(), "cpu_count": os.cpu_count(), "kind": "port",
            "workload": "cfg1: 4096 rays x 128 samples, render_weight_from_density + accumulate_along_rays "
                        "fwd+bwd, batched CPU tensors (the reference's torch.cumsum path restated)"}
