"""Fused scale / rotation / opacity activations between the deformation field and the rasterizer.

The reference applies `pc.scaling_activation` (torch.exp), `pc.rotation_activation` (F.normalize, dim=1,
eps=1e-12) and `pc.opacity_activation` (torch.sigmoid) to the deformation outputs
(gaussian_renderer/__init__.py:97-99; bound in scene/gaussian_model.py:47-58): 5 ATen kernels forward and
~10 in autograd.  `fused_activations` does the same arithmetic in one CUDA kernel each way through the C ABI
(`fdgs_activate_forward` / `fdgs_activate_backward`).  `render()` uses it only when the model's activation
functions ARE those three; anything else keeps the model's own callables.
"""
import ctypes as C

import torch

from . import _lib as L


class _FusedActivations(torch.autograd.Function):
    @staticmethod
    def forward(ctx, scales_raw, rotations_raw, opacity_raw):
        lib = L.load()
        s_raw, r_raw, o_raw = scales_raw.contiguous(), rotations_raw.contiguous(), opacity_raw.contiguous()
        s, r, o = torch.empty_like(s_raw), torch.empty_like(r_raw), torch.empty_like(o_raw)
        a = L.FdgsActivateForward(P=s_raw.shape[0], scales_raw=L.ptr(s_raw), rotations_raw=L.ptr(r_raw), opacity_raw=L.ptr(o_raw),
                                  scales=L.ptr(s), rotations=L.ptr(r), opacity=L.ptr(o))
        L.check(lib.fdgs_activate_forward(C.byref(a), L.stream_ptr()), "fdgs_activate_forward")
        ctx.save_for_backward(r_raw, s, r, o)
        for t, need in zip((s, r, o), ctx.needs_input_grad):
            if not need:
                ctx.mark_non_differentiable(t)
        return s, r, o

    @staticmethod
    def backward(ctx, g_s, g_r, g_o):
        lib = L.load()
        r_raw, s, r, o = ctx.saved_tensors
        need = ctx.needs_input_grad
        g_s = g_s.contiguous() if (g_s is not None and need[0]) else None
        g_r = g_r.contiguous() if (g_r is not None and need[1]) else None
        g_o = g_o.contiguous() if (g_o is not None and need[2]) else None
        d_s = torch.empty_like(s) if g_s is not None else None
        d_r = torch.empty_like(r) if g_r is not None else None
        d_o = torch.empty_like(o) if g_o is not None else None
        a = L.FdgsActivateBackward(P=s.shape[0], rotations_raw=L.ptr(r_raw), scales=L.ptr(s), rotations=L.ptr(r), opacity=L.ptr(o),
                                   g_scales=L.ptr(g_s), g_rotations=L.ptr(g_r), g_opacity=L.ptr(g_o),
                                   d_scales_raw=L.ptr(d_s), d_rotations_raw=L.ptr(d_r), d_opacity_raw=L.ptr(d_o))
        L.check(lib.fdgs_activate_backward(C.byref(a), L.stream_ptr()), "fdgs_activate_backward")
        return d_s, d_r, d_o


def fused_activations(scales_raw, rotations_raw, opacity_raw):
    """(exp(scales_raw), normalize(rotations_raw, dim=1), sigmoid(opacity_raw)) for [P,3], [P,4], [P,1] CUDA FP32 tensors."""
    for t, w in ((scales_raw, 3), (rotations_raw, 4), (opacity_raw, 1)):
        if not t.is_cuda or t.dtype != torch.float32 or t.dim() != 2 or t.shape[1] != w or t.shape[0] != scales_raw.shape[0]:
            raise ValueError("fused_activations expects CUDA float32 tensors of shape [P,3], [P,4], [P,1]")
    return _FusedActivations.apply(scales_raw, rotations_raw, opacity_raw)


def is_reference_activations(pc):
    """True when the model binds exactly the reference's three activation functions."""
    return (getattr(pc, "scaling_activation", None) is torch.exp and getattr(pc, "opacity_activation", None) is torch.sigmoid
            and getattr(pc, "rotation_activation", None) is torch.nn.functional.normalize)
