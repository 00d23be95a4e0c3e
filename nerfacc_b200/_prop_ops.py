"""Tensor-level wrappers and autograd Functions over the PropNetEstimator kernels of the C ABI
(nfb_cdf_from_density_fwd/bwd, nfb_pdf_loss_fwd/bwd; csrc/propnet.cu)."""
from __future__ import annotations

import torch
from torch import Tensor

from . import _C

_MAX_CDF_ROW = 512


def cdf_rows_ok(t_vals: Tensor, sigmas: Tensor) -> bool:
    """Batched float32 CUDA rows [n_rays, n+1] / [n_rays, n] the fused kernel takes."""
    return (t_vals.is_cuda and sigmas.is_cuda and t_vals.dim() == 2 and sigmas.dim() == 2
            and t_vals.dtype == torch.float32 and sigmas.dtype == torch.float32
            and t_vals.shape[0] == sigmas.shape[0] and t_vals.shape[1] == sigmas.shape[1] + 1
            and 1 <= sigmas.shape[1] <= _MAX_CDF_ROW)


class CdfFromDensity(torch.autograd.Function):
    """cdfs [R, n+1] = 1 - cat([exp(-exclusive_sum(sigmas * (t[1:] - t[:-1]))), 0]) — what
    PropNetEstimator.sampling builds per level (prop_net.py:100-114) — differentiable in sigmas."""

    @staticmethod
    def forward(ctx, t_vals: Tensor, sigmas: Tensor) -> Tensor:
        t_vals, sigmas = t_vals.contiguous(), sigmas.contiguous()
        dev = _C.require_cuda(t_vals, sigmas)
        n_rays, n = sigmas.shape
        cdfs = torch.empty_like(t_vals)
        with torch.cuda.device(dev):
            rc = _C.lib().nfb_cdf_from_density_fwd(t_vals.data_ptr(), sigmas.data_ptr(), n_rays, n,
                                                   cdfs.data_ptr(), _C.stream_ptr(dev))
        _C.check(rc, "cdf_from_density_fwd")
        ctx.save_for_backward(t_vals, sigmas)
        return cdfs

    @staticmethod
    def backward(ctx, g_cdfs: Tensor):
        t_vals, sigmas = ctx.saved_tensors
        if not ctx.needs_input_grad[1]:
            return None, None
        dev = sigmas.device
        n_rays, n = sigmas.shape
        g_cdfs = g_cdfs.contiguous()
        g_sigmas = torch.empty_like(sigmas)
        with torch.cuda.device(dev):
            rc = _C.lib().nfb_cdf_from_density_bwd(t_vals.data_ptr(), sigmas.data_ptr(), g_cdfs.data_ptr(),
                                                   n_rays, n, g_sigmas.data_ptr(), _C.stream_ptr(dev))
        _C.check(rc, "cdf_from_density_bwd")
        return None, g_sigmas


def pdf_loss_rows_ok(q_vals: Tensor, q_cdfs: Tensor, k_vals: Tensor, k_cdfs: Tensor) -> bool:
    ts = (q_vals, q_cdfs, k_vals, k_cdfs)
    return (all(t.is_cuda and t.dim() == 2 and t.dtype == torch.float32 for t in ts)
            and q_vals.shape == q_cdfs.shape and k_vals.shape == k_cdfs.shape
            and q_vals.shape[0] == k_vals.shape[0] and q_vals.shape[1] >= 2 and k_vals.shape[1] >= 1
            and 8 * 3 * k_vals.shape[1] * 4 <= 200 * 1024
            and not q_vals.requires_grad and not q_cdfs.requires_grad and not k_vals.requires_grad)


class PdfLoss(torch.autograd.Function):
    """_pdf_loss of prop_net.py:232-256 for batched rows, differentiable in cdfs_key (the only
    input that carries gradient in PropNetEstimator.compute_loss)."""

    @staticmethod
    def forward(ctx, q_vals: Tensor, q_cdfs: Tensor, k_vals: Tensor, k_cdfs: Tensor, eps: float) -> Tensor:
        q_vals, q_cdfs, k_vals, k_cdfs = (t.contiguous() for t in (q_vals, q_cdfs, k_vals, k_cdfs))
        dev = _C.require_cuda(q_vals, q_cdfs, k_vals, k_cdfs)
        n_rays, n_q = q_vals.shape
        n_k = k_vals.shape[1]
        loss = torch.empty((n_rays, n_q - 1), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            rc = _C.lib().nfb_pdf_loss_fwd(q_vals.data_ptr(), q_cdfs.data_ptr(), n_q, k_vals.data_ptr(),
                                           k_cdfs.data_ptr(), n_k, n_rays, float(eps), loss.data_ptr(),
                                           _C.stream_ptr(dev))
        _C.check(rc, "pdf_loss_fwd")
        ctx.save_for_backward(q_vals, q_cdfs, k_vals, k_cdfs)
        ctx.eps = float(eps)
        return loss

    @staticmethod
    def backward(ctx, g_loss: Tensor):
        q_vals, q_cdfs, k_vals, k_cdfs = ctx.saved_tensors
        if not ctx.needs_input_grad[3]:
            return None, None, None, None, None
        dev = k_cdfs.device
        n_rays, n_q = q_vals.shape
        n_k = k_vals.shape[1]
        g_loss = g_loss.contiguous()
        g_k = torch.empty_like(k_cdfs)
        with torch.cuda.device(dev):
            rc = _C.lib().nfb_pdf_loss_bwd(q_vals.data_ptr(), q_cdfs.data_ptr(), n_q, k_vals.data_ptr(),
                                           k_cdfs.data_ptr(), n_k, n_rays, ctx.eps, g_loss.data_ptr(),
                                           g_k.data_ptr(), _C.stream_ptr(dev))
        _C.check(rc, "pdf_loss_bwd")
        return None, None, None, g_k, None
