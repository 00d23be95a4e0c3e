"""ctypes binding of libnerfacc_b200.so (the C ABI declared in include/nerfacc_b200.h).

This module replaces the reference's lazy pybind shim (nerfacc/cuda/__init__.py:8-40 and
nerfacc/cuda/_backend.py).  There is no CPU fallback and no other backend: if the shared
library is missing or a tensor is not on a CUDA device, calls raise.

PyTorch is used for plumbing only: it owns every buffer (caching allocator), provides the
current stream, and its tensors' ``data_ptr()`` are what crosses the boundary.
"""
from __future__ import annotations

import ctypes as C
import os
import threading
from typing import Dict, Optional, Tuple

import torch

from . import _build

_LIB = None
_LOCK = threading.Lock()

c_f32p = C.c_void_p
c_i64 = C.c_int64
c_int = C.c_int
c_u32 = C.c_uint32
c_f32 = C.c_float
c_size = C.c_size_t
c_ptr = C.c_void_p



class Rays(C.Structure):
    """nfb_rays_t"""
    _fields_ = [("rays_o", c_ptr), ("rays_d", c_ptr), ("rays_mask", c_ptr), ("near_planes", c_ptr),
                ("far_planes", c_ptr), ("near_plane", c_f32), ("far_plane", c_f32),
                ("t_sorted", c_ptr), ("t_indices", c_ptr), ("hits", c_ptr), ("n_rays", c_i64)]


class Grid(C.Structure):
    """nfb_grid_t"""
    _fields_ = [("occ_bits", c_ptr), ("occ_mip", c_ptr), ("mip_shift", C.c_int32), ("aabbs", c_ptr),
                ("n_grids", C.c_int32), ("res", C.c_int32 * 3), ("occ_bounds", c_ptr)]


class March(C.Structure):
    """nfb_march_t"""
    _fields_ = [("step_size", c_f32), ("cone_angle", c_f32), ("traverse_steps_limit", C.c_int32)]


_sp = C.POINTER
# name -> argtypes; every function returns int unless listed in _RESTYPES.
_PROTOS = {
    "nfb_version": [],
    "nfb_compiled_arch": [],
    "nfb_last_error": [],
    "nfb_launch_count": [],
    "nfb_workspace_bytes": [c_i64],
    "nfb_scan_f32": [c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_int, c_int, c_int, c_int, c_int,
                     c_ptr, c_size, c_u32, c_ptr],
    "nfb_prod_scan_bwd_f32": [c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_int, c_int,
                              c_ptr, c_size, c_u32, c_ptr],
    "nfb_head_bits_bytes": [c_i64],
    "nfb_head_bits": [c_ptr, c_i64, c_i64, c_ptr, c_ptr],
    "nfb_scan_bits_f32": [c_ptr, c_ptr, c_ptr, c_i64, c_int, c_int, c_int, c_ptr, c_size, c_u32, c_ptr],
    "nfb_prod_scan_bwd_bits_f32": [c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_int, c_ptr, c_size, c_u32,
                                   c_ptr],
    "nfb_render_weight_fwd": [c_int, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_i64,
                              c_ptr, c_ptr, c_ptr, c_ptr, c_f32, c_f32, c_ptr, c_size, c_u32, c_ptr],
    "nfb_render_weight_bwd": [c_int, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr,
                              c_i64, c_i64, c_i64, c_ptr, c_ptr, c_size, c_u32, c_ptr],
    "nfb_render_weight_bits_fwd": [c_int, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_i64,
                                   c_ptr, c_ptr, c_ptr, c_ptr, c_f32, c_f32, c_ptr, c_size, c_u32, c_ptr],
    "nfb_render_weight_bits_bwd": [c_int, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr,
                                   c_i64, c_ptr, c_ptr, c_size, c_u32, c_ptr],
    "nfb_accumulate_fwd": [c_ptr, c_ptr, c_i64, c_ptr, c_ptr, c_i64, c_i64, c_ptr, c_int,
                           c_ptr, c_size, c_u32, c_ptr],
    "nfb_accumulate_bwd": [c_ptr, c_ptr, c_ptr, c_i64, c_ptr, c_i64, c_ptr, c_ptr, c_ptr],
    "nfb_composite_fwd": [c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_ptr, c_ptr, c_ptr,
                          c_ptr, c_size, c_u32, c_ptr],
    "nfb_composite_bwd": [c_int, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr,
                          c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_ptr, c_ptr, c_ptr, c_size, c_u32,
                          c_ptr],
    "nfb_render_rays_fwd": [c_int, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_ptr, c_ptr, c_ptr,
                            c_ptr, c_ptr, c_ptr, c_int, c_ptr],
    "nfb_render_rays_bwd": [c_int, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr,
                            c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_ptr, c_ptr, c_int, c_ptr],
    "nfb_render_bits_workspace_bytes": [c_i64],
    "nfb_render_bits_fwd": [c_int, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_ptr, c_ptr,
                            c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_size, c_u32, c_ptr],
    "nfb_render_bits_bwd": [c_int, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr,
                            c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_ptr, c_ptr, c_ptr, c_size, c_u32, c_ptr],
    "nfb_accumulate_bits_fwd": [c_ptr, c_ptr, c_i64, c_ptr, c_ptr, c_i64, c_i64, c_ptr, c_int, c_ptr, c_size, c_u32,
                                c_ptr],
    "nfb_depth_normalize_fwd": [c_ptr, c_ptr, c_i64, c_ptr],
    "nfb_depth_normalize_bwd": [c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_ptr, c_ptr, c_ptr],
    "nfb_pack_info": [c_ptr, c_i64, c_i64, c_ptr, c_ptr, c_size, c_ptr],
    "nfb_ray_aabb_intersect": [c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_f32, c_f32, c_f32, c_ptr, c_ptr,
                               c_ptr, c_ptr],
    "nfb_pack_occupancy": [c_ptr, C.c_int32, c_ptr, c_ptr, c_ptr, C.c_int32, c_ptr],
    "nfb_occupancy_bounds": [c_ptr, C.c_int32, c_ptr, c_ptr, c_ptr],
    "nfb_traverse_run_bytes": [c_i64],
    "nfb_traverse_count": [_sp(Rays), _sp(Grid), _sp(March), c_ptr, c_ptr, c_ptr, c_ptr, c_ptr,
                           c_size, c_ptr, c_size, c_ptr],
    "nfb_traverse_scratch_bytes": [c_i64],
    "nfb_traverse_expand": [c_i64, c_ptr, c_ptr, c_i64, c_ptr, c_ptr, c_ptr, c_ptr, c_size, c_ptr],
    "nfb_traverse_piece_bytes": [c_i64],
    "nfb_traverse_expand_cone": [c_i64, c_ptr, c_ptr, C.c_float, C.c_float, c_i64, c_ptr, c_ptr, c_ptr, c_ptr, c_size,
                                 c_ptr],
    "nfb_traverse_fill_sampling": [_sp(Rays), _sp(Grid), _sp(March), c_ptr, c_ptr, c_ptr, c_ptr, c_ptr,
                                   c_i64, c_ptr, c_size, c_ptr],
    "nfb_chunk_state_bytes": [],
    "nfb_chunk_init": [c_ptr, c_i64, C.c_int32, c_ptr, c_ptr],
    "nfb_chunk_march": [_sp(Rays), _sp(Grid), _sp(March), c_ptr, c_i64, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr,
                        c_ptr, c_ptr, c_ptr, c_ptr, c_size, c_ptr],
    "nfb_chunk_composite": [c_ptr, c_ptr, c_i64, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_ptr, c_ptr,
                            c_ptr, c_f32, c_f32, C.c_int32, c_ptr, c_ptr],
    "nfb_occ_cell_points": [c_ptr, c_ptr, c_i64, c_ptr, c_ptr, c_ptr, c_ptr],
    "nfb_occ_ema_update": [c_ptr, c_ptr, c_i64, c_ptr, c_i64, c_i64, c_f32, c_ptr, c_ptr],
    "nfb_occ_binarize_workspace_bytes": [],
    "nfb_occ_binarize": [c_ptr, c_i64, c_f32, c_ptr, c_ptr, c_ptr, c_size, c_ptr],
    "nfb_mark_invisible_cells": [c_ptr, c_i64, c_ptr, C.c_int32, c_ptr, c_f32, c_f32, c_f32, c_ptr, c_ptr],
    "nfb_occ_select_cells": [c_ptr, c_ptr, c_ptr, c_i64, c_ptr, c_ptr, c_ptr, c_ptr, c_size, c_ptr],
    "nfb_compact_workspace_bytes": [c_i64],
    "nfb_compact_samples": [c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr,
                            c_size, c_ptr],
    "nfb_visibility_compact": [c_int, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_ptr, c_f32, c_f32,
                               c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_size, c_u32, c_ptr],
    "nfb_compact_samples_static": [c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr],
    "nfb_importance_sampling": [c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_i64, c_ptr, c_i64, c_i64, c_ptr, c_ptr,
                                c_ptr, c_ptr, c_ptr, c_ptr],
    "nfb_sample_positions": [c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_ptr, c_ptr, c_ptr],
    "nfb_importance_sampling_stot": [c_ptr, c_ptr, c_i64, c_i64, c_i64, c_ptr, c_ptr, c_ptr, c_int, c_f32, c_f32,
                                     c_ptr],
    "nfb_cdf_from_density_fwd": [c_ptr, c_ptr, c_i64, c_i64, c_ptr, c_ptr],
    "nfb_cdf_from_density_bwd": [c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_ptr, c_ptr],
    "nfb_pdf_loss_fwd": [c_ptr, c_ptr, c_i64, c_ptr, c_ptr, c_i64, c_i64, c_f32, c_ptr, c_ptr],
    "nfb_pdf_loss_bwd": [c_ptr, c_ptr, c_i64, c_ptr, c_ptr, c_i64, c_i64, c_f32, c_ptr, c_ptr, c_ptr],
    "nfb_searchsorted": [c_ptr, c_ptr, c_i64, c_ptr, c_ptr, c_i64, c_i64, c_i64, c_ptr, c_ptr, c_ptr],
    "nfb_traverse_fill": [_sp(Rays), _sp(Grid), _sp(March), c_int, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr,
                          c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_size, c_ptr],
}
_RESTYPES = {
    "nfb_last_error": C.c_char_p,
    "nfb_launch_count": C.c_longlong,
    "nfb_workspace_bytes": C.c_size_t,
    "nfb_render_bits_workspace_bytes": C.c_size_t,
    "nfb_head_bits_bytes": C.c_size_t,
    "nfb_occ_binarize_workspace_bytes": C.c_size_t,
    "nfb_chunk_state_bytes": C.c_size_t,
    "nfb_compact_workspace_bytes": C.c_size_t,
    "nfb_traverse_run_bytes": C.c_size_t,
    "nfb_traverse_piece_bytes": C.c_size_t,
    "nfb_traverse_scratch_bytes": C.c_size_t,
}


def declared_symbols():
    return sorted(_PROTOS)


def lib():
    """Load (building first if the sources changed and nvcc is present) the CUDA library."""
    global _LIB
    if _LIB is not None:
        return _LIB
    with _LOCK:
        if _LIB is not None:
            return _LIB
        path = _build.LIB
        if os.environ.get("NFB_NO_BUILD", "0") != "1":
            try:
                if not _build.up_to_date():
                    _build.build()
            except Exception as e:  # no nvcc / read-only tree: use what is there
                if not os.path.exists(path):
                    raise RuntimeError(
                        "libnerfacc_b200.so is missing and could not be built "
                        f"({e}); nerfacc_b200 has no CPU or PyTorch fallback."
                    ) from e
        if not os.path.exists(path):
            raise RuntimeError(
                f"{path} not found; run `python -m nerfacc_b200._build`. "
                "nerfacc_b200 has no CPU or PyTorch fallback."
            )
        l = C.CDLL(path)
        proxy = _LibProxy(path)
        for name, argtypes in _PROTOS.items():
            fn = getattr(l, name)
            fn.argtypes = argtypes
            fn.restype = _RESTYPES.get(name, C.c_int)
            # entry points that launch kernels (they take the stream last) go through a
            # timer shim so bench.py can measure them live; queries are bound directly
            launches = name not in _RESTYPES and len(argtypes) > 0
            setattr(proxy, name, _Timed(name[4:], fn) if launches else fn)
        _LIB = proxy
    return _LIB


class _LibProxy:
    """Attribute bag holding the typed entry points of libnerfacc_b200.so."""

    def __init__(self, path: str) -> None:
        self._path = path


class _Timed:
    __slots__ = ("name", "fn")

    def __init__(self, name: str, fn) -> None:
        self.name, self.fn = name, fn

    def __call__(self, *args):
        if _PROF is None:
            return self.fn(*args)
        t0 = prof_start()
        rc = self.fn(*args)
        prof_stop(self.name, t0)
        return rc


def last_error() -> str:
    msg = lib().nfb_last_error()
    return msg.decode() if msg else ""


def launch_count() -> int:
    return int(lib().nfb_launch_count())


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise RuntimeError(f"nerfacc_b200.{what} failed (code {rc}): {last_error()}")


def require_cuda(*tensors: Optional[torch.Tensor]) -> torch.device:
    dev = None
    for t in tensors:
        if t is None:
            continue
        if not t.is_cuda:
            raise RuntimeError(
                "nerfacc_b200 only runs on CUDA tensors (no CPU fallback by design); "
                f"got a tensor on {t.device}"
            )
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise RuntimeError(f"tensors on different devices: {dev} vs {t.device}")
    assert dev is not None
    return dev


# --------------------------------------------------------------------------- kernel timers
# bench.py measures each native entry point live, inside its timed region, with CUDA events
# recorded on the launching stream right around the C-ABI call (torch.cuda.Event only sees
# torch's current stream, which is the stream the kernels are launched on).
_PROF: Optional[Dict[str, list]] = None


def prof_enable(on: bool = True) -> None:
    global _PROF
    _PROF = {} if on else None


def prof_start():
    if _PROF is None:
        return None
    e = torch.cuda.Event(enable_timing=True)
    e.record()
    return e


def prof_stop(name: str, start) -> None:
    if start is None or _PROF is None:
        return
    e = torch.cuda.Event(enable_timing=True)
    e.record()
    _PROF.setdefault(name, []).append((start, e))


def prof_collect() -> Dict[str, Tuple[int, float]]:
    """name -> (launches, total milliseconds); synchronises."""
    out: Dict[str, Tuple[int, float]] = {}
    if _PROF is None:
        return out
    torch.cuda.synchronize()
    for name, evs in _PROF.items():
        out[name] = (len(evs), float(sum(a.elapsed_time(b) for a, b in evs)))
    _PROF.clear()
    return out


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


_RAW_STREAM = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def stream_ptr(device: torch.device) -> int:
    """The current stream's handle on `device` (the raw accessor: a tenth of the cost of building a
    torch.cuda.Stream object, and this runs before every native call)."""
    if _RAW_STREAM is not None:
        return _RAW_STREAM(device.index if device.index is not None else torch.cuda.current_device())
    return torch.cuda.current_stream(device).cuda_stream


# --------------------------------------------------------------------------- workspace
class _Workspace:
    """Persistent look-back scratch per (device, stream): zero-filled once, then reused with
    a fresh epoch per launch so no memset is needed (see nfb_workspace_bytes)."""

    __slots__ = ("buf", "epoch")

    def __init__(self) -> None:
        self.buf: Optional[torch.Tensor] = None
        self.epoch = 0


_WS: Dict[Tuple[int, int], _Workspace] = {}
_EPOCH_MAX = (1 << 30) - 2


def workspace(device: torch.device, n_elements: int) -> Tuple[torch.Tensor, int, int]:
    """Returns (buffer, nbytes, epoch) for one flat-kernel launch on the current stream."""
    need = int(lib().nfb_workspace_bytes(max(int(n_elements), 1)))
    key = (device.index if device.index is not None else torch.cuda.current_device(), stream_ptr(device))
    ws = _WS.get(key)
    if ws is None:
        ws = _WS[key] = _Workspace()
    if torch.cuda.is_current_stream_capturing():
        # graph replays would re-use a baked epoch: let the call clear the scratch itself
        if ws.buf is None or ws.buf.numel() < need:
            raise RuntimeError("nerfacc_b200: warm up once outside CUDA-graph capture first")
        return ws.buf, ws.buf.numel(), 0
    if ws.buf is None or ws.buf.numel() < need:
        size = max(need * 2, 1 << 20)
        ws.buf = torch.zeros(size, dtype=torch.uint8, device=device)
        ws.epoch = 0
    ws.epoch += 1
    if ws.epoch > _EPOCH_MAX:
        ws.buf.zero_()
        ws.epoch = 1
    return ws.buf, ws.buf.numel(), ws.epoch


_SCRATCH: Dict[Tuple[int, int, str], torch.Tensor] = {}


def scratch(device: torch.device, nbytes: int, tag: str = "ray") -> torch.Tensor:
    """A persistent uint8 scratch buffer per (device, stream, tag), separate from the look-back
    workspace (kernels that clear their own scratch must not disturb its epochs)."""
    key = (device.index if device.index is not None else torch.cuda.current_device(),
           stream_ptr(device), tag)
    buf = _SCRATCH.get(key)
    if buf is None or buf.numel() < nbytes:
        if torch.cuda.is_current_stream_capturing() and buf is not None:
            raise RuntimeError("nerfacc_b200: warm up once outside CUDA-graph capture first")
        buf = torch.empty(max(int(nbytes) * 2, 1 << 16), dtype=torch.uint8, device=device)
        _SCRATCH[key] = buf
    return buf
