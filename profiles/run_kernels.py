"""One launch of every hot kernel at its benchmarked shape, for `ncu` (profiling tooling).

    ncu --set full --clock-control none --import-source on -o gpurun_out/r02_kernels \\
        python profiles/run_kernels.py [cfg2] [cfg3] [cfg4] [cfg5]
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench_configs as BC  # noqa: E402
import nerfacc_b200 as nf  # noqa: E402
from bench import AABB, RES, STEP_SIZE, make_grid, make_rays  # noqa: E402
from nerfacc_b200 import _occ_ops, _prop_ops  # noqa: E402

which = set(sys.argv[1:]) or {"cfg2", "cfg3", "cfg4", "cfg5"}
dev = torch.device("cuda:0")
g0 = torch.Generator(device=dev).manual_seed(1234)

if "cfg2" in which:
    n_rays = 1 << 20
    o_np, d_np = make_rays(n_rays, 1234)
    o, d = torch.from_numpy(o_np).to(dev), torch.from_numpy(d_np).to(dev)
    est = nf.OccGridEstimator(AABB.tolist(), resolution=RES, levels=1).to(dev)
    est.binaries = torch.from_numpy(make_grid()).to(dev)
    ri, ts, te = est.sampling(o, d, render_step_size=STEP_SIZE)
    S = ri.numel()
    sig = (torch.rand(S, device=dev, generator=g0) * 40).requires_grad_(True)
    rgb = torch.rand((S, 3), device=dev, generator=g0).requires_grad_(True)
    c, op, dp, _ = nf.rendering(ts, te, ri, n_rays, rgb_sigma_fn=lambda a, b, r: (rgb, sig))
    torch.autograd.backward([c, op, dp], [torch.rand_like(c), torch.rand_like(op), torch.rand_like(dp)])
    # sampling with a density callback: the one-pass visibility -> compaction
    est.occs = est.binaries.flatten().float()
    sig0 = sig.detach()
    est.sampling(o, d, sigma_fn=lambda a, b, r: sig0, render_step_size=STEP_SIZE)
    del ri, ts, te, sig, rgb, c, op, dp, sig0
    torch.cuda.empty_cache()

if "cfg3" in which:
    n_rays = 1 << 20
    cnt = torch.from_numpy(BC.cfg3_counts(n_rays, 1 << 24)).to(dev)
    ri = torch.repeat_interleave(torch.arange(n_rays, device=dev), cnt)
    pi = nf.pack_info(ri, n_rays)
    S = ri.numel()
    alphas = (torch.rand(S, device=dev, generator=g0) ** 4).requires_grad_(True)
    w, _ = nf.render_weight_from_alpha(alphas, packed_info=pi)
    torch.autograd.grad(w, alphas, torch.rand_like(w))
    ts = torch.rand(S, device=dev, generator=g0)
    _occ_ops.visibility_compact(False, alphas.detach(), ts, ts + 0.01, pi, ri, 1e-4, 1e-2)
    del ri, pi, alphas, w, ts
    torch.cuda.empty_cache()

if "cfg4" in which:
    n_rays = 1 << int(os.environ.get("NFB_CFG4_LOG2_RAYS", "17"))
    est = nf.OccGridEstimator(BC.CFG4["roi"], resolution=BC.CFG4["res"], levels=BC.CFG4["levels"]).to(dev)
    est.binaries = BC.cfg4_binaries(dev)
    o, d = BC.cfg4_rays(n_rays, 4321, dev)
    est.sampling(o, d, render_step_size=BC.CFG4["step"], cone_angle=BC.CFG4["cone"], near_plane=BC.CFG4["near"])
    est.train()
    est._update(0, BC.cfg4_occ_eval_fn)
    est._update(512, BC.cfg4_occ_eval_fn)
    torch.cuda.empty_cache()

if "cfg5" in which:
    n_rays = 1 << 20
    near, far = 0.2, 1e3
    cd = torch.tensor([0.0, 1.0], device=dev).repeat(n_rays, 1)
    lv = cd.clone()
    levels = []
    for n_out in (256, 96, 48):
        iv, _, t_vals = nf.importance_sampling(nf.RayIntervals(vals=lv), cd, n_out, _samples=False,
                                               _stot=("lindisp", near, far))
        lv = iv.vals
        if n_out != 48:
            sig = (torch.rand((n_rays, n_out), device=dev, generator=g0) ** 3 * 5).requires_grad_(True)
            cd = _prop_ops.CdfFromDensity.apply(t_vals, sig)
            torch.autograd.grad(cd, sig, torch.rand_like(cd))
            cd = cd.detach()
            levels.append((lv, cd))
    qc = torch.rand((n_rays, 49), device=dev).cumsum(-1)
    qc = qc / qc[:, -1:]
    kv, kc = levels[0]
    nf.searchsorted(nf.RayIntervals(vals=kv), nf.RayIntervals(vals=lv))
    kc = kc.clone().requires_grad_(True)
    loss = _prop_ops.PdfLoss.apply(lv, qc, kv, kc, 1e-7)
    torch.autograd.grad(loss, kc, torch.rand_like(loss))
torch.cuda.synchronize()
print("ran", sorted(which))
