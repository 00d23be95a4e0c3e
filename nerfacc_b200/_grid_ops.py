"""Tensor-level wrappers over the traversal entry points of the C ABI (nfb_traverse_*,
nfb_pack_occupancy, nfb_ray_aabb_intersect in include/nerfacc_b200.h)."""
from __future__ import annotations

import ctypes as C
import math
import weakref
from typing import Dict, Optional, Tuple

import torch
from torch import Tensor

from . import _C

_BITS: Dict[int, Tuple[weakref.ref, int, Tensor, Optional[Tensor], int, Optional[Tensor]]] = {}
_MIP_BYTES_MAX = 40 * 1024  # mip kept in shared memory by the marching kernels (2 CTAs per SM must fit)


def occupancy_bits(binaries: Tensor) -> Tuple[Tensor, Optional[Tensor], int, Optional[Tensor]]:
    """Bit-packed copy of a bool grid [m, rx, ry, rz] (int32 words, 32 cells each) and its
    "any occupied" mip (one bit per (2^shift)^3 block, sized to fit shared memory), cached per
    tensor object and in-place version so repeated marches through the same grid pack once.
    Returns (bits, mip or None, shift, bounds or None) — bounds: per level the index box of the
    occupied cells (nfb_occupancy_bounds)."""
    key = id(binaries)
    ent = _BITS.get(key)
    if ent is not None and ent[0]() is binaries and ent[1] == binaries._version:
        return ent[2], ent[3], ent[4], ent[5]
    dev = _C.require_cuda(binaries)
    b = binaries.contiguous()
    if b.dtype != torch.bool:
        b = b != 0
    m, res = b.shape[0], [int(v) for v in b.shape[1:]]
    n = b.numel()
    bits = torch.empty(((n + 31) // 32,), dtype=torch.int32, device=dev)
    mip, shift, bounds = None, 0, None
    if n:
        for sh in range(1, 9):
            words = (m * math.prod((r + (1 << sh) - 1) >> sh for r in res) + 31) // 32
            if words * 4 <= _MIP_BYTES_MAX:
                shift = sh
                mip = torch.empty((words,), dtype=torch.int32, device=dev)
                break
        res_host = (C.c_int32 * 3)(*res)
        with torch.cuda.device(dev):
            rc = _C.lib().nfb_pack_occupancy(b.data_ptr(), m, res_host, bits.data_ptr(),
                                             None if mip is None else mip.data_ptr(), shift,
                                             _C.stream_ptr(dev))
        _C.check(rc, "pack_occupancy")
        bounds = torch.empty((m, 6), dtype=torch.int32, device=dev)
        with torch.cuda.device(dev):
            rc = _C.lib().nfb_occupancy_bounds(bits.data_ptr(), m, res_host, bounds.data_ptr(),
                                               _C.stream_ptr(dev))
        _C.check(rc, "occupancy_bounds")

    def _drop(_r, key=key):
        _BITS.pop(key, None)

    _BITS[key] = (weakref.ref(binaries, _drop), binaries._version, bits, mip, shift, bounds)
    return bits, mip, shift, bounds


_PINNED: Dict[int, Tensor] = {}


def _pinned_totals(dev: torch.device) -> Tensor:
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    t = _PINNED.get(idx)
    if t is None:
        t = _PINNED[idx] = torch.zeros(4, dtype=torch.long).pin_memory()
    return t


class Marcher:
    """One traversal problem: keeps the tensors alive and owns the C descriptors."""

    def __init__(self, rays_o: Tensor, rays_d: Tensor, binaries: Tensor, aabbs: Tensor,
                 near_planes, far_planes, step_size: float, cone_angle: float, limit: int,
                 rays_mask: Optional[Tensor] = None, t_sorted: Optional[Tensor] = None,
                 t_indices: Optional[Tensor] = None, hits: Optional[Tensor] = None):
        self.dev = _C.require_cuda(rays_o, rays_d, binaries, aabbs)
        f32 = lambda t: t.contiguous() if t.dtype == torch.float32 else t.float().contiguous()
        self.rays_o, self.rays_d, self.aabbs = f32(rays_o), f32(rays_d), f32(aabbs)
        self.n_rays = self.rays_o.shape[0]
        assert binaries.dim() == 4 and binaries.shape[0] == self.aabbs.shape[0]
        self.bits, self.mip, mip_shift, self.bounds = occupancy_bits(binaries)
        keep = [self.rays_o, self.rays_d, self.aabbs, self.bits, self.mip, self.bounds]
        r = _C.Rays()
        r.rays_o, r.rays_d = self.rays_o.data_ptr(), self.rays_d.data_ptr()
        r.n_rays = self.n_rays
        for name, val, default in (("near", near_planes, 0.0), ("far", far_planes, float("inf"))):
            if isinstance(val, Tensor):
                t = f32(val)
                keep.append(t)
                setattr(r, name + "_planes", t.data_ptr())
            else:
                setattr(r, name + "_plane", float(default if val is None else val))
        if rays_mask is not None:
            m = rays_mask.contiguous()
            m = m if m.dtype == torch.bool else (m != 0)
            keep.append(m)
            r.rays_mask = m.data_ptr()
        if t_sorted is not None and t_indices is not None and hits is not None:
            ts, ti = f32(t_sorted), t_indices.long().contiguous()
            hh = hits.contiguous() if hits.dtype == torch.bool else (hits != 0).contiguous()
            keep += [ts, ti, hh]
            r.t_sorted, r.t_indices, r.hits = ts.data_ptr(), ti.data_ptr(), hh.data_ptr()
        g = _C.Grid()
        g.occ_bits, g.aabbs, g.n_grids = self.bits.data_ptr(), self.aabbs.data_ptr(), binaries.shape[0]
        if self.mip is not None:
            g.occ_mip, g.mip_shift = self.mip.data_ptr(), mip_shift
        if self.bounds is not None:
            g.occ_bounds = self.bounds.data_ptr()
        g.res[0], g.res[1], g.res[2] = binaries.shape[1], binaries.shape[2], binaries.shape[3]
        self._keep, self.rays, self.grid = keep, r, g
        self.march = _C.March(float(step_size), float(cone_angle), int(limit))
        self.run_buf: Optional[Tensor] = None
        self.n_overflowed = -1
        self._spec: Optional[Tuple[Tensor, Tensor, Tensor]] = None  # outputs expanded ahead of the read-back

    def _scratch(self) -> Tensor:
        return _C.scratch(self.dev, int(_C.lib().nfb_traverse_scratch_bytes(self.n_rays)), "ray")

    def count(self, want_intervals: bool, want_term: bool, want_runs: bool = False, expand_capacity: int = 0):
        """Pass 1.  Returns (sample_info, interval_info|None, termination|None, n_samples, n_edges);
        synchronises the stream once to learn the totals (the only host sync of a traversal).
        With `want_runs` (constant-dt marching only) the pass also records each ray's sample runs
        so that fill_sampling() can expand them instead of marching again.  With `expand_capacity`
        (the caller's guess of the total, e.g. the previous step's plus a margin) the runs are
        expanded into buffers of that many samples BEFORE the totals are read back, so the device
        works through the host round trip; fill_sampling() keeps them if the total fits."""
        dev, n = self.dev, self.n_rays
        sample_info = torch.empty((n, 2), dtype=torch.long, device=dev)
        interval_info = torch.empty((n, 2), dtype=torch.long, device=dev) if want_intervals else None
        term = torch.empty((n,), dtype=torch.float32, device=dev) if want_term else None
        if n == 0:
            return sample_info, interval_info, term, 0, 0
        totals = _pinned_totals(dev)
        self.run_buf = None
        with torch.cuda.device(dev):
            sc = self._scratch()
            cone = self.march.cone_angle != 0.0
            if want_runs and self.march.step_size > 0.0:
                # constant dt: runs (arithmetic progressions); cone stepping: the piece log
                size = _C.lib().nfb_traverse_piece_bytes(n) if cone else _C.lib().nfb_traverse_run_bytes(n)
                self.run_buf = _C.scratch(dev, int(size), "runs")
            rc = _C.lib().nfb_traverse_count(
                C.byref(self.rays), C.byref(self.grid), C.byref(self.march), sample_info.data_ptr(),
                None if interval_info is None else interval_info.data_ptr(),
                None if term is None else term.data_ptr(), totals.data_ptr(), sc.data_ptr(),
                sc.numel(), None if self.run_buf is None else self.run_buf.data_ptr(),
                0 if self.run_buf is None else self.run_buf.numel(), _C.stream_ptr(dev))
            _C.check(rc, "traverse_count")
            self._spec = None
            counted = None
            if self.run_buf is not None and expand_capacity > 0:
                # the host waits for the totals only (this event), not for the expansion queued behind them
                counted = torch.cuda.Event()
                counted.record(torch.cuda.current_stream(dev))
                cap = int(expand_capacity)
                spec = (torch.empty((cap,), dtype=torch.long, device=dev),
                        torch.empty((cap,), dtype=torch.float32, device=dev),
                        torch.empty((cap,), dtype=torch.float32, device=dev))
                if cone:
                    rc = _C.lib().nfb_traverse_expand_cone(
                        n, sample_info.data_ptr(), self.run_buf.data_ptr(), self.march.step_size, self.march.cone_angle,
                        cap, spec[0].data_ptr(), spec[1].data_ptr(), spec[2].data_ptr(), sc.data_ptr(), sc.numel(),
                        _C.stream_ptr(dev))
                    _C.check(rc, "traverse_expand_cone")
                else:
                    rc = _C.lib().nfb_traverse_expand(n, sample_info.data_ptr(), self.run_buf.data_ptr(), cap,
                                                      spec[0].data_ptr(), spec[1].data_ptr(), spec[2].data_ptr(),
                                                      sc.data_ptr(), sc.numel(), _C.stream_ptr(dev))
                    _C.check(rc, "traverse_expand")
                self._spec = spec
                counted.synchronize()
            else:
                torch.cuda.current_stream(dev).synchronize()
        self.n_overflowed = int(totals[2])
        if cone and self.run_buf is not None and self.n_overflowed != 0:
            self.run_buf, self._spec, self.n_overflowed = None, None, 0  # the piece log ran full: pass 2 marches
        return sample_info, interval_info, term, int(totals[0]), int(totals[1])

    def fill_sampling(self, sample_info: Tensor, n_samples: int):
        dev = self.dev
        spec, self._spec = self._spec, None
        if spec is not None and n_samples <= spec[0].shape[0]:
            # the runs were expanded while the totals travelled to the host
            ray_indices, t_starts, t_ends = (t[:n_samples] for t in spec)
            if n_samples > 0 and self.n_overflowed > 0 and self.march.cone_angle == 0.0:  # rays whose runs did not fit the buffer: march those
                with torch.cuda.device(dev):
                    sc = self._scratch()
                    rc = _C.lib().nfb_traverse_fill_sampling(
                        C.byref(self.rays), C.byref(self.grid), C.byref(self.march), sample_info.data_ptr(),
                        ray_indices.data_ptr(), t_starts.data_ptr(), t_ends.data_ptr(), self.run_buf.data_ptr(),
                        -self.n_overflowed - 1, sc.data_ptr(), sc.numel(), _C.stream_ptr(dev))
                _C.check(rc, "traverse_fill_sampling")
            return ray_indices, t_starts, t_ends
        ray_indices = torch.empty((n_samples,), dtype=torch.long, device=dev)
        t_starts = torch.empty((n_samples,), dtype=torch.float32, device=dev)
        t_ends = torch.empty((n_samples,), dtype=torch.float32, device=dev)
        if n_samples > 0:
            with torch.cuda.device(dev):
                sc = self._scratch()
                rc = _C.lib().nfb_traverse_fill_sampling(
                    C.byref(self.rays), C.byref(self.grid), C.byref(self.march),
                    sample_info.data_ptr(), ray_indices.data_ptr(), t_starts.data_ptr(),
                    t_ends.data_ptr(), None if self.run_buf is None else self.run_buf.data_ptr(),
                    self.n_overflowed, sc.data_ptr(), sc.numel(), _C.stream_ptr(dev))
            _C.check(rc, "traverse_fill_sampling")
        return ray_indices, t_starts, t_ends

    def fill_full(self, over_allocate: bool, sample_info: Tensor, interval_info: Tensor,
                  n_samples: int, n_edges: int, sample_starts: Optional[Tensor] = None,
                  interval_starts: Optional[Tensor] = None, term: Optional[Tensor] = None):
        dev = self.dev
        mk = torch.zeros if over_allocate else torch.empty
        iv = dict(vals=mk((n_edges,), dtype=torch.float32, device=dev),
                  ray_indices=mk((n_edges,), dtype=torch.long, device=dev),
                  is_left=mk((n_edges,), dtype=torch.bool, device=dev),
                  is_right=mk((n_edges,), dtype=torch.bool, device=dev))
        sm = dict(vals=mk((n_samples,), dtype=torch.float32, device=dev),
                  ray_indices=mk((n_samples,), dtype=torch.long, device=dev),
                  is_valid=mk((n_samples,), dtype=torch.bool, device=dev))
        if self.n_rays > 0 and (over_allocate or n_samples > 0):
            totals = _pinned_totals(dev)
            with torch.cuda.device(dev):
                sc = self._scratch()
                p = lambda t: None if t is None else t.data_ptr()
                rc = _C.lib().nfb_traverse_fill(
                    C.byref(self.rays), C.byref(self.grid), C.byref(self.march), int(over_allocate),
                    p(sample_starts), p(interval_starts), sample_info.data_ptr(),
                    interval_info.data_ptr(), p(term), totals.data_ptr(), iv["vals"].data_ptr(),
                    iv["ray_indices"].data_ptr(), iv["is_left"].data_ptr(), iv["is_right"].data_ptr(),
                    sm["vals"].data_ptr(), sm["ray_indices"].data_ptr(), sm["is_valid"].data_ptr(),
                    sc.data_ptr(), sc.numel(), _C.stream_ptr(dev))
            _C.check(rc, "traverse_fill")
        return iv, sm
