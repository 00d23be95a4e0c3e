"""Generate tests/golden/golden_cpu.npz by running the UNMODIFIED Python reference on CPU.

    python tests/golden/make_golden_cpu.py            (needs /root/reference; authoring box only)

Everything the reference can execute without a GPU: the batched scans and
volume-rendering composites (torch.cumsum / cumprod paths), accumulate_along_rays via
index_add_, autograd gradients of all of them, the pure-torch helpers `_ray_aabb_intersect`,
`_sample_from_weighted`, `_lossfun_outer`, `_transform_stot`, `_enlarge_aabb`, and
`OccGridEstimator._update` / `mark_invisible_cells` on CPU tensors.  Inputs and outputs are
stored together; tests/test_oracle_golden.py replays the inputs through the CPU oracle and
tests/test_gpu_golden.py through the CUDA library.
"""
import os
import sys

import numpy as np
import torch

REF = os.environ.get("NERFACC_REF", "/root/reference")
sys.path.insert(0, REF)
import nerfacc as ref  # noqa: E402
from nerfacc.estimators.prop_net import _lossfun_outer, _transform_stot  # noqa: E402
from nerfacc.grid import _enlarge_aabb, _ray_aabb_intersect  # noqa: E402
from nerfacc.pdf import _sample_from_weighted  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden_cpu.npz")
G = {}


def put(name, t):
    G[name] = t.detach().cpu().numpy() if isinstance(t, torch.Tensor) else np.asarray(t)


def main():
    torch.manual_seed(42)
    R, N = 24, 37

    # ---- batched scans (scan.py:42-44, 84-91, 133-135, 176-182) incl. gradients
    x = (torch.rand(R, N) * 0.9 + 0.05).requires_grad_(True)
    g = torch.rand(R, N)
    put("scan/x", x); put("scan/g", g)
    for name, fn in (("inclusive_sum", ref.inclusive_sum), ("exclusive_sum", ref.exclusive_sum),
                     ("inclusive_prod", ref.inclusive_prod), ("exclusive_prod", ref.exclusive_prod)):
        y = fn(x)
        (gx,) = torch.autograd.grad((y * g).sum(), x)
        put(f"scan/{name}", y); put(f"scan/{name}_grad", gx)

    # ---- batched volume rendering, forward + gradients (volrend.py:161-480)
    t = torch.sort(torch.rand(R, N + 1), dim=-1).values * 4.0
    ts, te = t[:, :-1].contiguous(), t[:, 1:].contiguous()
    sig = (torch.rand(R, N) * 6.0).requires_grad_(True)
    gw, gT, ga = torch.rand(R, N), torch.rand(R, N), torch.rand(R, N)
    prefix = torch.rand(R, 1) * 0.8 + 0.1
    put("vr/ts", ts); put("vr/te", te); put("vr/sig", sig); put("vr/gw", gw); put("vr/gT", gT)
    put("vr/ga", ga); put("vr/prefix", prefix)
    for tag, pt in (("", None), ("_prefix", prefix)):
        w, T, a = ref.render_weight_from_density(ts, te, sig, prefix_trans=pt)
        put(f"vr/density{tag}_w", w); put(f"vr/density{tag}_T", T); put(f"vr/density{tag}_a", a)
        if pt is not None:
            # the reference multiplies the prefix in place into an exp()/cumprod() output
            # (volrend.py:204-205, 262-263), which makes its own backward raise: forward only
            continue
        (gs,) = torch.autograd.grad((w * gw + T * gT + a * ga).sum(), sig)
        put(f"vr/density{tag}_grad", gs)
        (gs_w,) = torch.autograd.grad((ref.render_weight_from_density(ts, te, sig, prefix_trans=pt)[0] * gw).sum(), sig)
        put(f"vr/density{tag}_grad_wonly", gs_w)
    al = (torch.rand(R, N) ** 3).requires_grad_(True)
    put("vr/alphas", al)
    for tag, pt in (("", None), ("_prefix", prefix)):
        w, T = ref.render_weight_from_alpha(al, prefix_trans=pt)
        put(f"vr/alpha{tag}_w", w); put(f"vr/alpha{tag}_T", T)
        if pt is not None:
            continue
        (gal,) = torch.autograd.grad((w * gw + T * gT).sum(), al)
        put(f"vr/alpha{tag}_grad", gal)
    put("vr/vis_alpha", ref.render_visibility_from_alpha(al.detach(), early_stop_eps=0.05, alpha_thre=0.02))
    put("vr/vis_density", ref.render_visibility_from_density(ts, te, sig.detach(), early_stop_eps=0.03, alpha_thre=0.1))

    # ---- accumulate_along_rays with ray_indices on CPU (index_add_, volrend.py:532-547)
    counts = torch.tensor([0, 5, 1, 0, 0, 17, 3, 64, 0, 2])
    ri = torch.repeat_interleave(torch.arange(len(counts)), counts)
    S = int(counts.sum())
    wv = torch.rand(S, requires_grad=True)
    vv = torch.rand(S, 3, requires_grad=True)
    go = torch.rand(len(counts), 3)
    out = ref.accumulate_along_rays(wv, vv, ri, len(counts))
    g_w, g_v = torch.autograd.grad((out * go).sum(), (wv, vv))
    put("acc/counts", counts); put("acc/w", wv); put("acc/v", vv); put("acc/go", go)
    put("acc/out", out); put("acc/gw", g_w); put("acc/gv", g_v)
    put("acc/out_none", ref.accumulate_along_rays(wv.detach(), None, ri, len(counts)))

    # ---- rendering() batched (volrend.py:14-158)
    rgb = torch.rand(R, N, 3)
    bk = torch.tensor([0.2, 0.5, 0.9])
    c, o, d, _ = ref.rendering(ts, te, rgb_sigma_fn=lambda a, b, r: (rgb, sig.detach()), render_bkgd=bk)
    put("rend/rgb", rgb); put("rend/bkgd", bk); put("rend/colors", c); put("rend/opac", o); put("rend/depth", d)

    # ---- slab test (grid.py:54-90) and aabb helper (grid.py:195-198)
    ro = torch.rand(200, 3) * 2 - 1
    rd = torch.randn(200, 3); rd = rd / rd.norm(dim=-1, keepdim=True)
    lo = torch.rand(7, 3) - 1.0
    bb = torch.cat([lo, lo + torch.rand(7, 3) + 0.1], -1)
    tmin, tmax, hits = _ray_aabb_intersect(ro, rd, bb)
    put("aabb/o", ro); put("aabb/d", rd); put("aabb/bb", bb); put("aabb/tmin", tmin); put("aabb/tmax", tmax)
    put("aabb/hits", hits)
    put("aabb/enlarged", _enlarge_aabb(torch.tensor([-1.0, -0.5, 0.0, 1.0, 2.0, 0.25]), 4.0))

    # ---- inverse-CDF sampling torch reference (pdf.py:134-219), as in tests/test_pdf.py:66-94
    B, NS = 100, 32
    bins = torch.sort(torch.rand(R, B + 1), dim=-1).values
    weights = torch.rand(R, B)
    # per ray, with the ray's own first/last edge as clamp (as tests/test_pdf.py:84-92 does)
    ec = [_sample_from_weighted(bins[i:i + 1], weights[i:i + 1], NS, False, bins[i].min(), bins[i].max())
          for i in range(R)]
    edges, centers = torch.cat([e for e, _ in ec]), torch.cat([c for _, c in ec])
    pdf = weights / weights.sum(-1, keepdim=True)
    cdfs = torch.cat([torch.zeros(R, 1), torch.cumsum(pdf, -1)], -1)
    put("pdf/bins", bins); put("pdf/cdfs", cdfs); put("pdf/edges", edges); put("pdf/centers", centers)

    # ---- proposal loss torch reference (prop_net.py:259-313), as in tests/test_pdf.py:98-127
    t_q = torch.sort(torch.rand(R, 33), dim=-1).values
    w_q = torch.rand(R, 32)
    t_k = torch.sort(torch.rand(R, 65), dim=-1).values
    w_k = torch.rand(R, 64)
    put("loss/t_q", t_q); put("loss/w_q", w_q); put("loss/t_k", t_k); put("loss/w_k", w_k)
    put("loss/outer", _lossfun_outer(t_q, w_q, t_k, w_k))
    s = torch.rand(R, 9)
    put("stot/s", s)
    put("stot/lindisp", _transform_stot("lindisp", s, torch.tensor(0.2), torch.tensor(1e3)))
    put("stot/uniform", _transform_stot("uniform", s, torch.tensor(0.2), torch.tensor(1e3)))

    # ---- OccGridEstimator on CPU: mark_invisible_cells + one warm-up _update (occ_grid.py:263-404)
    est = ref.OccGridEstimator([-1.0, -1.0, -1.0, 1.0, 1.0, 1.0], resolution=16, levels=2)
    K = torch.tensor([[[20.0, 0, 10.0], [0, 20.0, 10.0], [0, 0, 1]]])
    pose = torch.tensor([[[-1.0, 0.0, 0.0, 0.0], [0.0, 1.0, 0.0, 0.0], [0.0, 0.0, -1.0, 2.5]]])
    est.mark_invisible_cells(K, pose, 20, 20)
    put("occ/after_mark", est.occs)

    def occ_fn(x):  # constant inside each finest-level cell -> scatter order is immaterial
        cell = torch.floor((x + 2.0) / 4.0 * 32).clamp(0, 31)
        h = (cell[:, 0] * 7 + cell[:, 1] * 13 + cell[:, 2] * 29) % 17
        return (h[:, None] / 17.0 * 0.05).float()

    xs, occs_seen = [], []

    def spy(x):
        xs.append(x.clone())
        v = occ_fn(x)
        occs_seen.append(v.clone())
        return v

    torch.manual_seed(7)
    est._update(step=0, occ_eval_fn=spy, occ_thre=0.01, ema_decay=0.95, warmup_steps=256)
    cells = est._get_all_cells()  # unchanged by _update (occs >= 0 stays >= 0)
    torch.manual_seed(7)
    for lvl in range(2):
        put(f"occ/cells{lvl}", cells[lvl])
        put(f"occ/u{lvl}", torch.rand((cells[lvl].shape[0], 3)))
        put(f"occ/x{lvl}", xs[lvl]); put(f"occ/occ{lvl}", occs_seen[lvl])
    put("occ/aabbs", est.aabbs); put("occ/after_update", est.occs); put("occ/binaries", est.binaries)

    np.savez_compressed(OUT, **G)
    print(f"wrote {OUT}: {len(G)} arrays, {os.path.getsize(OUT) / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
