"""Fused L1 photometric loss (utils/loss_utils.py:20-21 `l1_loss(network_output, gt)` = mean |a - b|, train.py:244).

ATen runs sub, abs, mean (+ the batch division) forward and five more kernels in autograd over the full image;
here it is one sweep each way through the C ABI (`fdgs_l1_loss_forward` / `fdgs_l1_loss_backward`).
"""
import ctypes as C

import torch

from . import _lib as L


class _L1Loss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, b, scale):
        lib = L.load()
        a_c, b_c = a.contiguous(), b.contiguous()
        loss = torch.zeros((), dtype=torch.float32, device=a.device)
        L.check(lib.fdgs_l1_loss_forward(L.ptr(a_c), L.ptr(b_c), a_c.numel(), float(scale), loss.data_ptr(),
                                         L.stream_ptr()), "fdgs_l1_loss_forward")
        ctx.save_for_backward(a_c, b_c)
        ctx.scale = float(scale)
        return loss

    @staticmethod
    def backward(ctx, g):
        lib = L.load()
        a_c, b_c = ctx.saved_tensors
        g = g.contiguous().to(torch.float32)
        d_a = torch.empty_like(a_c)
        L.check(lib.fdgs_l1_loss_backward(L.ptr(a_c), L.ptr(b_c), a_c.numel(), ctx.scale, g.data_ptr(), d_a.data_ptr(),
                                          L.stream_ptr()), "fdgs_l1_loss_backward")
        d_b = -d_a if ctx.needs_input_grad[1] else None
        return d_a, d_b, None


def l1_loss(network_output, gt, scale=None):
    """mean |network_output - gt| (same shape, CUDA float32); `scale` overrides the 1/numel factor."""
    if network_output.shape != gt.shape:
        raise ValueError("l1_loss: shapes differ (the fused kernel does not broadcast)")
    if not network_output.is_cuda or network_output.dtype != torch.float32 or gt.dtype != torch.float32:
        raise ValueError("l1_loss expects CUDA float32 tensors")
    n = network_output.numel()
    if n == 0:
        return (network_output - gt).abs().mean()
    return _L1Loss.apply(network_output, gt, (1.0 / n) if scale is None else scale)


def l1_view_loss(image, gt, n_views_total):
    """Fused counterpart of fdgs.dp.l1_view_loss: this view's share mean|img - gt| / B of the batch loss."""
    return l1_loss(image, gt, scale=1.0 / (image.numel() * float(n_views_total)))
