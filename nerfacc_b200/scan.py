"""Packed / batched prefix scans — drop-in for nerfacc/scan.py (same names, signatures,
asserts and autograd behaviour; reference nerfacc/scan.py:12-288).

Batched (N-D) inputs go through torch.cumsum / torch.cumprod exactly as in the reference
(scan.py:42-44, 84-91, 133-135, 176-182).  Packed (flattened) inputs go through the C ABI of
libnerfacc_b200: `nfb_scan_f32` / `nfb_prod_scan_bwd_f32` (include/nerfacc_b200.h).
"""
from typing import Optional

import torch
from torch import Tensor

from . import _C
from ._packing import as_packed_info, head_bits

_OP_SUM, _OP_PROD = 0, 1


def _scan(inputs: Tensor, packed_info: Tensor, canonical: bool, op: int, exclusive: bool,
          reverse: bool, normalize: bool = False) -> Tensor:
    dev = _C.require_cuda(inputs, packed_info)
    if inputs.dtype != torch.float32:
        raise RuntimeError(f"packed scans support float32 only, got {inputs.dtype}")
    outputs = torch.empty_like(inputs)
    n = inputs.numel()
    if n == 0 or packed_info.shape[0] == 0:
        return outputs
    bits = head_bits(packed_info, n) if (canonical and not normalize) else None
    with torch.cuda.device(dev):
        ws, ws_bytes, epoch = _C.workspace(dev, n)
        if bits is not None:
            rc = _C.lib().nfb_scan_bits_f32(
                inputs.data_ptr(), outputs.data_ptr(), bits.data_ptr(), n, op, int(exclusive),
                int(reverse), ws.data_ptr(), ws_bytes, epoch, _C.stream_ptr(dev))
            _C.check(rc, "scan")
            return outputs
        rc = _C.lib().nfb_scan_f32(
            inputs.data_ptr(), outputs.data_ptr(), packed_info.data_ptr(), packed_info.shape[0],
            n, op, int(exclusive), int(reverse), int(normalize), int(canonical),
            ws.data_ptr(), ws_bytes, epoch, _C.stream_ptr(dev))
    _C.check(rc, "scan")
    return outputs


def _prod_bwd(grad_outputs: Tensor, inputs: Tensor, outputs: Tensor, packed_info: Tensor,
              canonical: bool, exclusive: bool) -> Tensor:
    dev = _C.require_cuda(grad_outputs, inputs, outputs, packed_info)
    grad_inputs = torch.empty_like(grad_outputs)
    n = inputs.numel()
    if n == 0 or packed_info.shape[0] == 0:
        return grad_inputs
    bits = head_bits(packed_info, n) if canonical else None
    with torch.cuda.device(dev):
        ws, ws_bytes, epoch = _C.workspace(dev, n)
        if bits is not None:
            rc = _C.lib().nfb_prod_scan_bwd_bits_f32(
                grad_outputs.data_ptr(), inputs.data_ptr(), outputs.data_ptr(), grad_inputs.data_ptr(),
                bits.data_ptr(), n, int(exclusive), ws.data_ptr(), ws_bytes, epoch, _C.stream_ptr(dev))
            _C.check(rc, "prod_scan_bwd")
            return grad_inputs
        rc = _C.lib().nfb_prod_scan_bwd_f32(
            grad_outputs.data_ptr(), inputs.data_ptr(), outputs.data_ptr(), grad_inputs.data_ptr(),
            packed_info.data_ptr(), packed_info.shape[0], n, int(exclusive), int(canonical),
            ws.data_ptr(), ws_bytes, epoch, _C.stream_ptr(dev))
    _C.check(rc, "prod_scan_bwd")
    return grad_inputs


def inclusive_sum(inputs: Tensor, packed_info: Optional[Tensor] = None) -> Tensor:
    """Inclusive sum along the last dim, or per chunk of a flattened tensor
    (`packed_info` (n_rays, 2) = start, count).  Reference: nerfacc/scan.py:12-54.

    >>> inputs = torch.tensor([1., 2., 3., 4., 5., 6., 7., 8., 9.], device="cuda")
    >>> packed_info = torch.tensor([[0, 2], [2, 3], [5, 4]], device="cuda")
    >>> inclusive_sum(inputs, packed_info)
    tensor([ 1.,  3.,  3.,  7., 12.,  6., 13., 21., 30.], device='cuda:0')
    """
    if packed_info is None:
        outputs = torch.cumsum(inputs, dim=-1)
    else:
        assert inputs.dim() == 1, "inputs must be flattened."
        packed_info, canon = as_packed_info(packed_info)
        outputs = _InclusiveSum.apply(packed_info, canon, inputs, False)
    return outputs


def exclusive_sum(inputs: Tensor, packed_info: Optional[Tensor] = None) -> Tensor:
    """Exclusive sum; see :func:`inclusive_sum`.  Reference: nerfacc/scan.py:57-100."""
    if packed_info is None:
        outputs = torch.cumsum(
            torch.cat([torch.zeros_like(inputs[..., :1]), inputs[..., :-1]], dim=-1), dim=-1
        )
    else:
        assert inputs.dim() == 1, "inputs must be flattened."
        packed_info, canon = as_packed_info(packed_info)
        outputs = _ExclusiveSum.apply(packed_info, canon, inputs, False)
    return outputs


def inclusive_prod(inputs: Tensor, packed_info: Optional[Tensor] = None) -> Tensor:
    """Inclusive product; see :func:`inclusive_sum`.  Reference: nerfacc/scan.py:103-144."""
    if packed_info is None:
        outputs = torch.cumprod(inputs, dim=-1)
    else:
        assert inputs.dim() == 1, "inputs must be flattened."
        packed_info, canon = as_packed_info(packed_info)
        outputs = _InclusiveProd.apply(packed_info, canon, inputs)
    return outputs


def exclusive_prod(inputs: Tensor, packed_info: Optional[Tensor] = None) -> Tensor:
    """Exclusive product; see :func:`inclusive_sum`.  Reference: nerfacc/scan.py:147-186."""
    if packed_info is None:
        outputs = torch.cumprod(
            torch.cat([torch.ones_like(inputs[..., :1]), inputs[..., :-1]], dim=-1), dim=-1
        )
    else:
        packed_info, canon = as_packed_info(packed_info)
        outputs = _ExclusiveProd.apply(packed_info, canon, inputs)
    return outputs


class _InclusiveSum(torch.autograd.Function):
    """Inclusive Sum on a Flattened Tensor (reference nerfacc/scan.py:189-214)."""

    @staticmethod
    def forward(ctx, packed_info, canonical: bool, inputs, normalize: bool = False):
        inputs = inputs.contiguous()
        outputs = _scan(inputs, packed_info, canonical, _OP_SUM, False, False, normalize)
        if ctx.needs_input_grad[2]:
            ctx.normalize = normalize
            ctx.canonical = canonical
            ctx.save_for_backward(packed_info)
        return outputs

    @staticmethod
    def backward(ctx, grad_outputs):
        grad_outputs = grad_outputs.contiguous()
        (packed_info,) = ctx.saved_tensors
        assert ctx.normalize == False, "Only support backward for normalize==False."
        grad_inputs = _scan(grad_outputs, packed_info, ctx.canonical, _OP_SUM, False, True)
        return None, None, grad_inputs, None


class _ExclusiveSum(torch.autograd.Function):
    """Exclusive Sum on a Flattened Tensor (reference nerfacc/scan.py:217-242)."""

    @staticmethod
    def forward(ctx, packed_info, canonical: bool, inputs, normalize: bool = False):
        inputs = inputs.contiguous()
        outputs = _scan(inputs, packed_info, canonical, _OP_SUM, True, False, normalize)
        if ctx.needs_input_grad[2]:
            ctx.normalize = normalize
            ctx.canonical = canonical
            ctx.save_for_backward(packed_info)
        return outputs

    @staticmethod
    def backward(ctx, grad_outputs):
        grad_outputs = grad_outputs.contiguous()
        (packed_info,) = ctx.saved_tensors
        assert ctx.normalize == False, "Only support backward for normalize==False."
        grad_inputs = _scan(grad_outputs, packed_info, ctx.canonical, _OP_SUM, True, True)
        return None, None, grad_inputs, None


class _InclusiveProd(torch.autograd.Function):
    """Inclusive Product on a Flattened Tensor (reference nerfacc/scan.py:245-265)."""

    @staticmethod
    def forward(ctx, packed_info, canonical: bool, inputs):
        inputs = inputs.contiguous()
        outputs = _scan(inputs, packed_info, canonical, _OP_PROD, False, False)
        if ctx.needs_input_grad[2]:
            ctx.canonical = canonical
            ctx.save_for_backward(packed_info, inputs, outputs)
        return outputs

    @staticmethod
    def backward(ctx, grad_outputs):
        grad_outputs = grad_outputs.contiguous()
        packed_info, inputs, outputs = ctx.saved_tensors
        grad_inputs = _prod_bwd(grad_outputs, inputs, outputs, packed_info, ctx.canonical, False)
        return None, None, grad_inputs


class _ExclusiveProd(torch.autograd.Function):
    """Exclusive Product on a Flattened Tensor (reference nerfacc/scan.py:268-288)."""

    @staticmethod
    def forward(ctx, packed_info, canonical: bool, inputs):
        inputs = inputs.contiguous()
        outputs = _scan(inputs, packed_info, canonical, _OP_PROD, True, False)
        if ctx.needs_input_grad[2]:
            ctx.canonical = canonical
            ctx.save_for_backward(packed_info, inputs, outputs)
        return outputs

    @staticmethod
    def backward(ctx, grad_outputs):
        grad_outputs = grad_outputs.contiguous()
        packed_info, inputs, outputs = ctx.saved_tensors
        grad_inputs = _prod_bwd(grad_outputs, inputs, outputs, packed_info, ctx.canonical, True)
        return None, None, grad_inputs
