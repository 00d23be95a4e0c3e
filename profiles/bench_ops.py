"""Per-op micro-benchmarks on one B200: this library next to the reference's own CUDA kernels
(oracle/_ref/nerfacc_cuda.so, unmodified, sm_100a build) on identical device tensors — the
like-for-like bar of SURVEY.md §8(d).  Measurement tooling, not a product path.

    python profiles/bench_ops.py [--out gpurun_out/ops.json] [--quick]

Method: CUDA events on the current stream, 5 warm-ups, then 7 timed groups of 10 back-to-back
calls (no host sync inside a group, as in a training loop that keeps the stream busy; a
sync-bracketed single call adds 25-35 us of Python launch latency to these 60-200 us kernels),
median over groups of the per-call time; inputs are larger than L2 (126 MB) for every S-sized
op, so consecutive calls do not find their data cached.  GB/s = algorithmic bytes
/ time.  Both libraries are timed the same way.
"""
import argparse
import json
import os
import statistics
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import nerfacc_b200 as nf  # noqa: E402
from bench import AABB, RES, STEP_SIZE, make_grid, make_rays  # noqa: E402
from tests import _refcuda as R  # noqa: E402

PEAK = 6555.8
try:
    with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
        PEAK = float(json.load(f).get("hbm_gbs", PEAK))
except Exception:
    pass


def timeit(fn, warm=5, groups=7, reps=10):
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


ROWS = []


def row(name, algo_bytes, ours, ref=None, note=""):
    t_o = timeit(ours)
    t_r = timeit(ref) if ref is not None else None
    gbs = algo_bytes / (t_o * 1e-3) / 1e9
    ent = {"op": name, "bytes": algo_bytes, "ours_ms": t_o, "ours_gbs": gbs, "ours_frac": gbs / PEAK,
           "ref_ms": t_r, "speedup": (t_r / t_o) if t_r else None, "note": note}
    ROWS.append(ent)
    print(f"{name:44s} ours {t_o:8.3f} ms {gbs:7.0f} GB/s {100 * gbs / PEAK:5.1f}%   "
          f"ref {'' if t_r is None else f'{t_r:8.3f} ms  x{t_r / t_o:5.2f}'}  {note}", flush=True)


class _RefExclusiveSum(torch.autograd.Function):
    """The reference's _ExclusiveSum (nerfacc/scan.py:218-242) on its own kernels."""

    @staticmethod
    def forward(ctx, x, pi):
        ctx.pi = pi
        return R.scan(x, pi, "sum", True)

    @staticmethod
    def backward(ctx, g):
        return R.scan(g.contiguous(), ctx.pi, "sum", True, backward=True), None


def ref_weight_from_density(ts, te, sig, pi):
    """nerfacc/volrend.py:209-264 + 312-362 composed as the reference does."""
    sd = sig * (te - ts)
    alphas = 1.0 - torch.exp(-sd)
    trans = torch.exp(-_RefExclusiveSum.apply(sd, pi))
    return trans * alphas, trans, alphas


def ref_accumulate(w, v, ri, n_rays):
    """nerfacc/volrend.py:532-547."""
    src = w[:, None] * v if v is not None else w[:, None]
    out = torch.zeros((n_rays, src.shape[-1]), device=w.device, dtype=src.dtype)
    out.index_add_(0, ri, src)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "ops.json"))
    ap.add_argument("--quick", action="store_true", help="2^18 rays instead of 2^20")
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    have_ref = R.available()
    n_rays = 1 << (18 if args.quick else 20)

    # ------------------------------------------------------------ cfg2: marching
    o_np, d_np = make_rays(n_rays, 1234)
    o, d = torch.from_numpy(o_np).to(dev), torch.from_numpy(d_np).to(dev)
    bins = torch.from_numpy(make_grid()).to(dev)
    est = nf.OccGridEstimator(AABB.tolist(), resolution=RES, levels=1).to(dev)
    est.binaries = bins
    ri, ts, te = est.sampling(o, d, render_step_size=STEP_SIZE)
    S = ri.numel()
    pi = nf.pack_info(ri, n_rays)
    near = torch.zeros(n_rays, device=dev)
    far = torch.full((n_rays,), 1e10, device=dev)

    def ref_sampling():
        iv, sm, _ = R.traverse_grids(o, d, bins, est.aabbs, near, far, STEP_SIZE, 0.0)
        return sm["ray_indices"], iv["vals"][iv["is_left"]], iv["vals"][iv["is_right"]]

    row(f"OccGridEstimator.sampling cfg2 R=2^{n_rays.bit_length() - 1} S={S / 1e6:.1f}M", 16 * S + 40 * n_rays + 2 ** 18,
        lambda: est.sampling(o, d, render_step_size=STEP_SIZE), ref_sampling if have_ref else None,
        "ref = ray_aabb + sort + traverse_grids + is_left/is_right gathers (occ_grid.py:163-190)")

    # ------------------------------------------------------------ cfg3-like: scans on the marched layout
    g0 = torch.Generator(device=dev).manual_seed(1234)
    x = torch.rand(S, device=dev, generator=g0) * 0.1 + 0.9
    g = torch.rand(S, device=dev, generator=g0)
    for name, fn, op, excl in (("inclusive_sum", nf.inclusive_sum, "sum", False),
                               ("exclusive_sum", nf.exclusive_sum, "sum", True),
                               ("inclusive_prod", nf.inclusive_prod, "prod", False),
                               ("exclusive_prod", nf.exclusive_prod, "prod", True)):
        row(f"{name} fwd (packed)", 8 * S + 16 * n_rays, lambda fn=fn: fn(x, pi),
            (lambda op=op, excl=excl: R.scan(x, pi, op, excl)) if have_ref else None)
    y = nf.exclusive_prod(x, pi)
    from nerfacc_b200.scan import _prod_bwd
    row("exclusive_prod bwd (packed)", 16 * S + 16 * n_rays, lambda: _prod_bwd(g, x, y, pi, True, True),
        (lambda: R.prod_backward(g, x, y, pi, True)) if have_ref else None)

    # ------------------------------------------------------------ weights fwd / fwd+bwd
    sig = torch.rand(S, device=dev, generator=g0) * 10
    rgb = torch.rand((S, 3), device=dev, generator=g0)
    with torch.no_grad():
        row("render_weight_from_density fwd", 24 * S + 16 * n_rays,
            lambda: nf.render_weight_from_density(ts, te, sig, packed_info=pi),
            (lambda: ref_weight_from_density(ts, te, sig, pi)) if have_ref else None,
            "ref = volrend.py:209-264,312-362 on the reference's exclusive_sum kernel")
    sg = sig.clone().requires_grad_(True)

    def ours_wfb():
        w, _, _ = nf.render_weight_from_density(ts, te, sg, packed_info=pi)
        (gs,) = torch.autograd.grad(w, sg, g)
        return gs

    def ref_wfb():
        w, _, _ = ref_weight_from_density(ts, te, sg, pi)
        (gs,) = torch.autograd.grad(w, sg, g)
        return gs

    row("render_weight_from_density fwd+bwd", 48 * S + 32 * n_rays, ours_wfb, ref_wfb if have_ref else None)

    # ------------------------------------------------------------ accumulate
    w = torch.rand(S, device=dev, generator=g0)
    with torch.no_grad():
        row("accumulate_along_rays D=3", 24 * S + 28 * n_rays, lambda: nf.accumulate_along_rays(w, rgb, ri, n_rays),
            lambda: ref_accumulate(w, rgb, ri, n_rays), "ref = torch index_add_ (volrend.py:532-547)")
        row("accumulate_along_rays D=1 (values=None)", 12 * S + 20 * n_rays,
            lambda: nf.accumulate_along_rays(w, None, ri, n_rays), lambda: ref_accumulate(w, None, ri, n_rays))
        row("pack_info", 8 * S + 16 * n_rays, lambda: nf.pack_info(ri.clone(), n_rays), None,
            "includes an 8*S clone (cache defeat)")

    # ------------------------------------------------------------ rendering fwd+bwd (the bench step's second half)
    rg = rgb.clone().requires_grad_(True)

    def ours_render():
        c, op, dp, _ = nf.rendering(ts, te, ri, n_rays, rgb_sigma_fn=lambda a, b, r: (rg, sg))
        gs, gr = torch.autograd.grad((c.sum() + op.sum() + dp.sum()), (sg, rg))
        return gs

    def ref_pack_info():  # nerfacc/pack.py:38-46, run at every render_* call of the reference
        cnts = torch.zeros((n_rays,), device=dev, dtype=torch.long)
        cnts.index_add_(0, ri, torch.ones_like(ri))
        return torch.stack([cnts.cumsum(0) - cnts, cnts], -1)

    def ref_render():
        wts, _, _ = ref_weight_from_density(ts, te, sg, ref_pack_info())
        c = ref_accumulate(wts, rg, ri, n_rays)
        op = ref_accumulate(wts, None, ri, n_rays)
        dp = ref_accumulate(wts, ((ts + te) * 0.5)[:, None], ri, n_rays)
        dp = dp / op.clamp_min(torch.finfo(rg.dtype).eps)
        gs, gr = torch.autograd.grad((c.sum() + op.sum() + dp.sum()), (sg, rg))
        return gs

    row("rendering fwd+bwd (rgb_sigma_fn = stored tensors)", (24 + 24 + 52) * S + (32 + 72) * n_rays, ours_render,
        ref_render if have_ref else None, "ref = volrend.py:14-158 composed on the reference's kernels")
    del x, g, y, sig, rgb, w, sg, rg
    torch.cuda.empty_cache()

    # ------------------------------------------------------------ cfg5: proposal resampling 2 -> 257 -> 97 -> 49 edges
    n5 = n_rays
    cd = torch.tensor([0.0, 1.0], device=dev).repeat(n5, 1)
    lv = cd.clone()
    for n_out in (256, 96, 48):
        n_in = lv.shape[1]
        row(f"importance_sampling [{n5},{n_in}] -> n={n_out}", 8 * n_in * n5 + (8 * n_out + 4) * n5,
            lambda lv=lv, cd=cd, n_out=n_out: nf.importance_sampling(nf.RayIntervals(vals=lv), cd, n_out),
            (lambda lv=lv, cd=cd, n_out=n_out: R.importance_sampling(lv, cd, n_out)) if have_ref else None)
        row(f"  ... intervals only (PropNetEstimator's call)", 8 * n_in * n5 + (4 * n_out + 8) * n5,
            lambda lv=lv, cd=cd, n_out=n_out: nf.importance_sampling(nf.RayIntervals(vals=lv), cd, n_out, _samples=False))
        iv, _ = nf.importance_sampling(nf.RayIntervals(vals=lv), cd, n_out)
        lv = iv.vals
        wgt = torch.rand((n5, n_out), device=dev, generator=g0) + 1e-3
        cd = torch.cat([torch.zeros((n5, 1), device=dev), torch.cumsum(wgt, -1)], -1)
        cd = cd / cd[:, -1:]
        if n_out == 256:
            key = lv
    q = lv  # [n5, 49] final edges searched in the first level's 257 edges (the proposal loss)
    row(f"searchsorted query [{n5},{q.shape[1]}] in keys [{n5},{key.shape[1]}]",
        4 * (q.shape[1] + key.shape[1]) * n5 + 16 * q.shape[1] * n5,
        lambda: nf.searchsorted(nf.RayIntervals(vals=key), nf.RayIntervals(vals=q)),
        (lambda: R.searchsorted(key, q)) if have_ref else None)
    del cd, lv, key, q
    torch.cuda.empty_cache()

    # ------------------------------------------------------------ cfg4: 4 nested levels, cone stepping
    n4 = n_rays >> 3  # 2^17 rays per GPU at 8 GPUs
    o4, d4 = o[:n4].contiguous(), d[:n4].contiguous()
    est4 = nf.OccGridEstimator([-0.75, -0.75, -0.75, 0.75, 0.75, 0.75], resolution=RES, levels=4).to(dev)
    b4 = torch.rand((4, RES, RES, RES), device=dev, generator=g0) < 0.05
    est4.binaries = b4
    ri4, _, _ = est4.sampling(o4, d4, render_step_size=STEP_SIZE, cone_angle=0.004, near_plane=0.05)
    S4 = ri4.numel()
    near4 = torch.full((n4,), 0.05, device=dev)
    far4 = torch.full((n4,), 1e10, device=dev)

    def ref_sampling4():
        iv, sm, _ = R.traverse_grids(o4, d4, b4, est4.aabbs, near4, far4, STEP_SIZE, 0.004)
        return sm["ray_indices"], iv["vals"][iv["is_left"]], iv["vals"][iv["is_right"]]

    row(f"OccGridEstimator.sampling cfg4 m=4 cone R=2^{n4.bit_length() - 1} S={S4 / 1e6:.1f}M", 16 * S4 + 40 * n4 + 2 ** 20,
        lambda: est4.sampling(o4, d4, render_step_size=STEP_SIZE, cone_angle=0.004, near_plane=0.05),
        ref_sampling4 if have_ref else None)

    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    with open(args.out, "w") as f:
        json.dump({"peak_gbs": PEAK, "n_rays": n_rays, "samples": S, "device": torch.cuda.get_device_name(0),
                   "rows": ROWS}, f, indent=1)
    print("\n| op | algorithmic MB | ours ms | GB/s | of peak | reference CUDA ms | speed-up |")
    print("|---|---|---|---|---|---|---|")
    for r in ROWS:
        ref = "" if r["ref_ms"] is None else f"{r['ref_ms']:.3f}"
        sp = "" if r["speedup"] is None else f"{r['speedup']:.1f}x"
        print(f"| {r['op']} | {r['bytes'] / 1e6:.0f} | {r['ours_ms']:.3f} | {r['ours_gbs']:.0f} | "
              f"{100 * r['ours_frac']:.0f}% | {ref} | {sp} |")


if __name__ == "__main__":
    main()
