"""CPU oracle for the nerfacc hot path — numpy front-end of oracle/nfb_oracle.c.

TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this package; the product
(nerfacc_b200/) never does and fails loudly without its CUDA library instead.

Every function takes and returns numpy arrays (float32 / int64 / bool) and mirrors
one reference entry point; the arithmetic lives in C, each C function citing the
reference file:line it restates.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import numpy as np

from . import build as _build

_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(_build.build())
        _lib.orc_occ_binarize.restype = C.c_float
        _lib.orc_num_threads.restype = C.c_int
    return _lib


def num_threads() -> int:
    return int(lib().orc_num_threads())


def set_num_threads(n: int) -> None:
    lib().orc_set_num_threads(C.c_int(int(n)))


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _i64(a):
    return np.ascontiguousarray(a, dtype=np.int64)


def _u8(a):
    return np.ascontiguousarray(np.asarray(a).astype(np.uint8))


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


_i64c, _i32c, _f32c = C.c_int64, C.c_int32, C.c_float


# ----------------------------------------------------------------------------- scans
def scan(inputs, packed_info, op: str = "sum", exclusive: bool = False, reverse: bool = False,
         normalize: bool = False) -> np.ndarray:
    x, pi = _f32(inputs), _i64(packed_info)
    out = np.zeros_like(x)
    lib().orc_scan_f32(_p(x), _p(out), _p(pi), _i64c(pi.shape[0]), C.c_int(op == "prod"),
                       C.c_int(exclusive), C.c_int(reverse), C.c_int(normalize))
    return out


def scan_f64(inputs, packed_info, op: str = "sum", exclusive: bool = False,
             reverse: bool = False) -> np.ndarray:
    x, pi = _f32(inputs), _i64(packed_info)
    out = np.zeros(x.shape, dtype=np.float64)
    lib().orc_scan_f64(_p(x), _p(out), _p(pi), _i64c(pi.shape[0]), C.c_int(op == "prod"),
                       C.c_int(exclusive), C.c_int(reverse))
    return out


def inclusive_sum(x, pi, backward=False):
    return scan(x, pi, "sum", False, backward)


def exclusive_sum(x, pi, backward=False):
    return scan(x, pi, "sum", True, backward)


def inclusive_prod(x, pi):
    return scan(x, pi, "prod", False)


def exclusive_prod(x, pi):
    return scan(x, pi, "prod", True)


def prod_scan_backward(grad_out, inputs, outputs, packed_info, exclusive: bool) -> np.ndarray:
    g, x, y, pi = _f32(grad_out), _f32(inputs), _f32(outputs), _i64(packed_info)
    gi = np.zeros_like(x)
    lib().orc_prod_scan_bwd_f32(_p(g), _p(x), _p(y), _p(gi), _p(pi), _i64c(pi.shape[0]),
                                _i64c(x.shape[0]), C.c_int(exclusive))
    return gi


# ----------------------------------------------------------------------------- packing
def pack_info(ray_indices, n_rays: int) -> np.ndarray:
    ri = _i64(ray_indices)
    out = np.zeros((n_rays, 2), dtype=np.int64)
    lib().orc_pack_info(_p(ri), _i64c(ri.shape[0]), _i64c(n_rays), _p(out))
    return out


# ----------------------------------------------------------------------------- volrend
def render_weight_from_alpha(alphas, packed_info, prefix_trans=None):
    a, pi = _f32(alphas), _i64(packed_info)
    pt = None if prefix_trans is None else _f32(prefix_trans)
    w, T = np.zeros_like(a), np.zeros_like(a)
    lib().orc_weight_from_alpha(_p(a), _p(pt), _p(pi), _i64c(pi.shape[0]), _i64c(a.shape[0]),
                                _p(w), _p(T))
    return w, T


def render_weight_from_alpha_bwd(g_w, g_T, alphas, packed_info, prefix_trans=None):
    a, pi = _f32(alphas), _i64(packed_info)
    gw = None if g_w is None else _f32(g_w)
    gT = None if g_T is None else _f32(g_T)
    pt = None if prefix_trans is None else _f32(prefix_trans)
    ga = np.zeros_like(a)
    lib().orc_weight_from_alpha_bwd(_p(gw), _p(gT), _p(a), _p(pt), _p(pi), _i64c(pi.shape[0]),
                                    _i64c(a.shape[0]), _p(ga))
    return ga


def render_weight_from_density(t_starts, t_ends, sigmas, packed_info, prefix_trans=None):
    ts, te, s, pi = _f32(t_starts), _f32(t_ends), _f32(sigmas), _i64(packed_info)
    pt = None if prefix_trans is None else _f32(prefix_trans)
    w, T, a = np.zeros_like(s), np.zeros_like(s), np.zeros_like(s)
    lib().orc_weight_from_density(_p(ts), _p(te), _p(s), _p(pt), _p(pi), _i64c(pi.shape[0]),
                                  _i64c(s.shape[0]), _p(w), _p(T), _p(a))
    return w, T, a


def render_weight_from_density_bwd(g_w, g_T, g_a, t_starts, t_ends, sigmas, packed_info,
                                   prefix_trans=None):
    ts, te, s, pi = _f32(t_starts), _f32(t_ends), _f32(sigmas), _i64(packed_info)
    gw = None if g_w is None else _f32(g_w)
    gT = None if g_T is None else _f32(g_T)
    ga = None if g_a is None else _f32(g_a)
    pt = None if prefix_trans is None else _f32(prefix_trans)
    gs = np.zeros_like(s)
    lib().orc_weight_from_density_bwd(_p(gw), _p(gT), _p(ga), _p(ts), _p(te), _p(s), _p(pt),
                                      _p(pi), _i64c(pi.shape[0]), _i64c(s.shape[0]), _p(gs))
    return gs


def visibility(trans, alphas, early_stop_eps: float, alpha_thre: float) -> np.ndarray:
    T, a = _f32(trans), _f32(alphas)
    vis = np.zeros(T.shape, dtype=np.uint8)
    lib().orc_visibility(_p(T), _p(a), _i64c(T.shape[0]), _f32c(early_stop_eps),
                         _f32c(alpha_thre), _p(vis))
    return vis.astype(bool)


def accumulate_along_rays(weights, values, ray_indices, n_rays: int, out=None) -> np.ndarray:
    w, ri = _f32(weights), _i64(ray_indices)
    v = None if values is None else _f32(values)
    dim = 1 if v is None else v.shape[-1]
    if out is None:
        out = np.zeros((n_rays, dim), dtype=np.float32)
    assert out.dtype == np.float32 and out.flags.c_contiguous
    lib().orc_accumulate(_p(w), _p(v), _i64c(dim), _p(ri), _i64c(w.shape[0]), _i64c(n_rays), _p(out))
    return out


def accumulate_along_rays_bwd(g_out, weights, values, ray_indices):
    g, w, ri = _f32(g_out), _f32(weights), _i64(ray_indices)
    v = None if values is None else _f32(values)
    dim = g.shape[-1]
    gw = np.zeros_like(w)
    gv = None if v is None else np.zeros_like(v)
    lib().orc_accumulate_bwd(_p(g), _p(w), _p(v), _i64c(dim), _p(ri), _i64c(w.shape[0]), _p(gw), _p(gv))
    return gw, gv


# ----------------------------------------------------------------------------- grid
def ray_aabb_intersect(rays_o, rays_d, aabbs, near_plane=-np.inf, far_plane=np.inf,
                       miss_value=np.inf):
    o, d, bb = _f32(rays_o), _f32(rays_d), _f32(aabbs)
    n, m = o.shape[0], bb.shape[0]
    tmin, tmax = np.zeros((n, m), np.float32), np.zeros((n, m), np.float32)
    hits = np.zeros((n, m), np.uint8)
    lib().orc_ray_aabb_intersect(_p(o), _p(d), _p(bb), _i64c(n), _i64c(m), _f32c(near_plane),
                                 _f32c(far_plane), _f32c(miss_value), _p(tmin), _p(tmax), _p(hits))
    return tmin, tmax, hits.astype(bool)


def sort_intersections(t_mins, t_maxs):
    """torch.sort(cat([t_mins, t_maxs], -1), -1) of nerfacc/grid.py:160-162 (stable on ties)."""
    t = np.concatenate([t_mins, t_maxs], axis=-1)
    idx = np.argsort(t, axis=-1, kind="stable").astype(np.int64)
    return np.take_along_axis(t, idx, axis=-1), idx


def traverse_grids(rays_o, rays_d, binaries, aabbs, near_planes=None, far_planes=None,
                   step_size: float = 1e-3, cone_angle: float = 0.0,
                   traverse_steps_limit: Optional[int] = None, over_allocate: bool = False,
                   rays_mask=None, t_sorted=None, t_indices=None, hits=None) -> dict:
    """nerfacc/grid.py:94-192 + csrc/grid.cu:320-474.  Returns a dict of numpy arrays:
    intervals {vals, ray_indices, is_left, is_right, packed_info}, samples {vals,
    ray_indices, is_valid, packed_info}, termination_planes."""
    o, d, bb = _f32(rays_o), _f32(rays_d), _f32(aabbs)
    bins = _u8(binaries)
    n, m = o.shape[0], bins.shape[0]
    res = np.asarray(bins.shape[1:], dtype=np.int32)
    near = np.zeros(n, np.float32) if near_planes is None else _f32(near_planes)
    far = np.full(n, np.inf, np.float32) if far_planes is None else _f32(far_planes)
    limit = -1 if traverse_steps_limit is None else int(traverse_steps_limit)
    if over_allocate:
        assert limit > 0
    if t_sorted is None or t_indices is None or hits is None:
        tmin, tmax, hits = ray_aabb_intersect(o, d, bb)
        t_sorted, t_indices = sort_intersections(tmin, tmax)
    ts, ti, hh = _f32(t_sorted), _i64(t_indices), _u8(hits)
    # rays_mask is honoured only in the over-allocate path (grid.cu:380 vs :418,450)
    mask = None
    if over_allocate:
        mask = np.ones(n, np.uint8) if rays_mask is None else _u8(rays_mask)
    cnt_iv, cnt_sm = np.zeros(n, np.int64), np.zeros(n, np.int64)
    term = np.zeros(n, np.float32)
    L = lib()
    L.orc_traverse_count(_p(o), _p(d), _p(mask), _i64c(n), _i32c(m), _p(res), _p(bins), _p(bb),
                         _p(hh), _p(ts), _p(ti), _p(near), _p(far), _f32c(step_size),
                         _f32c(cone_angle), _i32c(limit), _p(cnt_iv), _p(cnt_sm), _p(term))
    if over_allocate:
        alloc_iv = mask.astype(np.int64) * (2 * limit)
        alloc_sm = mask.astype(np.int64) * limit
    else:
        alloc_iv, alloc_sm = cnt_iv, cnt_sm
    st_iv = np.cumsum(alloc_iv) - alloc_iv
    st_sm = np.cumsum(alloc_sm) - alloc_sm
    E, S = int(alloc_iv.sum()), int(alloc_sm.sum())
    iv_vals, iv_ray = np.zeros(E, np.float32), np.zeros(E, np.int64)
    iv_l, iv_r = np.zeros(E, np.uint8), np.zeros(E, np.uint8)
    sm_vals, sm_ray, sm_valid = np.zeros(S, np.float32), np.zeros(S, np.int64), np.zeros(S, np.uint8)
    L.orc_traverse_fill(_p(o), _p(d), _p(mask), _i64c(n), _i32c(m), _p(res), _p(bins), _p(bb),
                        _p(hh), _p(ts), _p(ti), _p(near), _p(far), _f32c(step_size),
                        _f32c(cone_angle), _i32c(limit), _p(st_iv), _p(st_sm), _p(iv_vals),
                        _p(iv_ray), _p(iv_l), _p(iv_r), _p(sm_vals), _p(sm_ray), _p(sm_valid))
    # packed_info describes the compact layout (grid.cu:402-404)
    cst_iv = np.cumsum(cnt_iv) - cnt_iv
    cst_sm = np.cumsum(cnt_sm) - cnt_sm
    return dict(
        intervals=dict(vals=iv_vals, ray_indices=iv_ray, is_left=iv_l.astype(bool),
                       is_right=iv_r.astype(bool), packed_info=np.stack([cst_iv, cnt_iv], -1)),
        samples=dict(vals=sm_vals, ray_indices=sm_ray, is_valid=sm_valid.astype(bool),
                     packed_info=np.stack([cst_sm, cnt_sm], -1)),
        termination_planes=term,
    )


def sampling(rays_o, rays_d, binaries, aabbs, near_plane=0.0, far_plane=1e10, t_min=None,
             t_max=None, render_step_size=1e-3, cone_angle=0.0) -> Tuple[np.ndarray, ...]:
    """OccGridEstimator.sampling without sigma_fn/alpha_fn (estimators/occ_grid.py:154-178)."""
    o = _f32(rays_o)
    near = np.full(o.shape[0], near_plane, np.float32)
    far = np.full(o.shape[0], far_plane, np.float32)
    if t_min is not None:
        near = np.maximum(near, _f32(t_min))
    if t_max is not None:
        far = np.minimum(far, _f32(t_max))
    r = traverse_grids(o, rays_d, binaries, aabbs, near, far, render_step_size, cone_angle)
    iv, sm = r["intervals"], r["samples"]
    return sm["ray_indices"], iv["vals"][iv["is_left"]], iv["vals"][iv["is_right"]], sm["packed_info"]


# ----------------------------------------------------------------------------- pdf
def importance_sampling(vals, cdfs, n_out: int, packed_info=None, bias=None):
    """Returns (intervals.vals [R, n_out+1], samples.vals [R, n_out]); csrc/pdf.cu:359-421."""
    v, c = _f32(vals), _f32(cdfs)
    b = None if bias is None else _f32(bias)
    if packed_info is None:
        assert v.ndim == 2
        n_rays, n_in = v.shape
        edges = np.zeros((n_rays, n_out + 1), np.float32)
        cen = np.zeros((n_rays, n_out), np.float32)
        lib().orc_importance_sampling_batched(_p(v), _p(c), _i64c(n_rays), _i64c(n_in),
                                              _i64c(n_out), _p(b), _p(edges), _p(cen))
    else:
        pi = _i64(packed_info)
        n_rays = pi.shape[0]
        edges = np.zeros((n_rays, n_out + 1), np.float32)
        cen = np.zeros((n_rays, n_out), np.float32)
        lib().orc_importance_sampling_packed(_p(v), _p(c), _p(pi), _i64c(n_rays), _i64c(n_out),
                                             _p(b), _p(edges), _p(cen))
    return edges, cen


def searchsorted(key, query, key_info=None, query_info=None):
    """(ids_left, ids_right) of nerfacc.searchsorted(sorted_sequence=key, values=query)."""
    k, q = _f32(key), _f32(query)
    il, ir = np.zeros(q.shape, np.int64), np.zeros(q.shape, np.int64)
    if key_info is None:
        assert k.ndim == 2 and q.ndim == 2
        lib().orc_searchsorted_batched(_p(q), _p(k), _i64c(q.shape[0]), _i64c(q.shape[1]),
                                       _i64c(k.shape[1]), _p(il), _p(ir))
    else:
        ki, qi = _i64(key_info), _i64(query_info)
        lib().orc_searchsorted_packed(_p(q), _p(qi), _p(k), _p(ki), _i64c(ki.shape[0]), _p(il), _p(ir))
    return il, ir


# ----------------------------------------------------------------------------- occupancy grid
def occ_cell_points(cells, u, resolution, aabb) -> np.ndarray:
    cl, uu = _i64(cells), _f32(u)
    res, bb = np.asarray(resolution, np.int32), _f32(aabb)
    x = np.zeros((cl.shape[0], 3), np.float32)
    lib().orc_occ_cell_points(_p(cl), _p(uu), _i64c(cl.shape[0]), _p(res), _p(bb), _p(x))
    return x


def occ_ema_update(occs, cell_ids, occ, decay: float) -> np.ndarray:
    o = _f32(occs).copy()
    ci, oc = _i64(cell_ids), _f32(occ)
    lib().orc_occ_ema_update(_p(o), _p(ci), _p(oc), _i64c(ci.shape[0]), _f32c(decay), _i64c(o.shape[0]))
    return o


def occ_binarize(occs, occ_thre: float):
    o = _f32(occs)
    b = np.zeros(o.shape, np.uint8)
    thre = lib().orc_occ_binarize(_p(o), _i64c(o.shape[0]), _f32c(occ_thre), _p(b))
    return b.astype(bool), float(thre)
