"""Loader + thin callers for the UNMODIFIED reference CUDA extension built by
oracle/build_ref.py into oracle/_ref/nerfacc_cuda.so (test infrastructure; GPU only).

The reference's Python package does not travel to the GPU box, so the pybind entry points
(registered at nerfacc/cuda/csrc/nerfacc.cpp:100-129) are called directly with the argument
conventions of the reference's own callers (nerfacc/scan.py:197-285, grid.py:43,165,
pdf.py:59,125, data_specs.py:47-60, 137-156).
"""
import importlib.util
import os

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "oracle", "_ref", "nerfacc_cuda.so")
_mod = None


def available() -> bool:
    return os.path.exists(SO) and torch.cuda.is_available()


def mod():
    global _mod
    if _mod is None:
        spec = importlib.util.spec_from_file_location("nerfacc_cuda", SO)
        _mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(_mod)
    return _mod


def _starts_cnts(packed_info):
    return packed_info[:, 0].contiguous(), packed_info[:, 1].contiguous()


def scan(x, packed_info, op="sum", exclusive=False, backward=False):
    s, c = _starts_cnts(packed_info)
    m = mod()
    if op == "sum":
        fn = m.exclusive_sum if exclusive else m.inclusive_sum
        return fn(s, c, x.contiguous(), False, backward)
    fn = m.exclusive_prod_forward if exclusive else m.inclusive_prod_forward
    return fn(s, c, x.contiguous())


def prod_backward(g, x, y, packed_info, exclusive):
    s, c = _starts_cnts(packed_info)
    fn = mod().exclusive_prod_backward if exclusive else mod().inclusive_prod_backward
    return fn(s, c, x.contiguous(), y.contiguous(), g.contiguous())


def ray_aabb_intersect(o, d, bb, near=-float("inf"), far=float("inf"), miss=float("inf")):
    return mod().ray_aabb_intersect(o.contiguous(), d.contiguous(), bb.contiguous(), near, far, miss)


def traverse_grids(o, d, binaries, aabbs, near=None, far=None, step_size=1e-3, cone_angle=0.0,
                   limit=None, over_allocate=False, rays_mask=None):
    """Call pattern of nerfacc/grid.py:94-192."""
    n = o.shape[0]
    if near is None:
        near = torch.zeros(n, device=o.device)
    if far is None:
        far = torch.full((n,), float("inf"), device=o.device)
    if rays_mask is None:
        rays_mask = torch.ones(n, dtype=torch.bool, device=o.device)
    t_mins, t_maxs, hits = ray_aabb_intersect(o, d, aabbs)
    t_sorted, t_indices = torch.sort(torch.cat([t_mins, t_maxs], -1), -1)
    iv, sm, term = mod().traverse_grids(
        o.contiguous(), d.contiguous(), rays_mask.contiguous(), binaries.contiguous(), aabbs.contiguous(),
        t_sorted.contiguous(), t_indices.contiguous(), hits.contiguous(), near.contiguous(),
        far.contiguous(), float(step_size), float(cone_angle), True, True, True,
        -1 if limit is None else int(limit), bool(over_allocate))
    unpack = lambda s, keys: {k: getattr(s, k) for k in keys} | {
        "packed_info": torch.stack([s.chunk_starts, s.chunk_cnts], -1)}
    return (unpack(iv, ("vals", "ray_indices", "is_left", "is_right")),
            unpack(sm, ("vals", "ray_indices", "is_valid")), term)


def _spec(vals, packed_info=None):
    s = mod().RaySegmentsSpec()
    s.vals = vals.contiguous()
    if packed_info is not None:
        s.chunk_starts, s.chunk_cnts = _starts_cnts(packed_info)
    return s


def importance_sampling(vals, cdfs, n, stratified=False):
    iv, sm = mod().importance_sampling(_spec(vals), cdfs.contiguous(), int(n), bool(stratified))
    return iv.vals, sm.vals


def searchsorted(key, query, key_info=None, query_info=None):
    return mod().searchsorted(_spec(query, query_info), _spec(key, key_info))
