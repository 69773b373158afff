"""HexPlane regularisers of the training step, fused.

Mirrors  scene/regulation.py:22-28        compute_plane_smoothness
         scene/gaussian_model.py:538-577  _plane_regulation / _time_regulation / _l1_regulation / compute_regulation
(train.py:251-254 adds `compute_regulation(time_smoothness_weight, l1_time_planes, plane_tv_weight)` to the loss in the
fine stage).  The reference evaluates them with ~40 ATen kernels forward and as many in autograd over the 6 x L plane
tensors; here ONE kernel (`fdgs_plane_regulation`, csrc/train_glue.cu) produces the value and the gradient of all of it.

Two ways in:
  * `compute_regulation(grids, ...)` and the three partial sums are differentiable functions with the reference's
    semantics (any upstream gradient; the plane gradients the kernel produced are scaled by it in backward);
  * `add_regulation_(grids, ...)` is the fully fused training path: it returns the value and ADDS the gradient straight
    into the planes' existing `.grad` tensors (e.g. the slots of the all-reduce bucket, where the deformation backward
    already wrote the sampler's plane gradients) — valid when the regulariser enters the loss with weight 1, which is
    how train.py uses it.  View-independent: in the view-parallel setting call it on ONE rank before the all-reduce, or
    on every rank after it.
"""
import ctypes as C

import torch

from . import _lib as L

SPATIAL, SPACETIME = (0, 1, 3), (2, 4, 5)


def _grids_of(obj):
    """Accepts the ModuleList `deformation_net.grid.grids`, a deform_network / Deformation, or a list of 6-plane lists."""
    for path in (("deformation_net", "grid", "grids"), ("grid", "grids"), ("grids",)):
        o = obj
        try:
            for a in path:
                o = getattr(o, a)
            return [list(level) for level in o]
        except AttributeError:
            continue
    return [list(level) for level in obj]


def _fill(a, grids, weights, grad_ptrs):
    levels = len(grids)
    if levels > L.FDGS_MAX_LEVELS:
        raise ValueError("too many HexPlane levels")
    a.levels = levels
    a.channels = int(grids[0][0].shape[1])
    a.plane_tv_weight, a.time_smoothness_weight, a.l1_time_planes_weight = [float(w) for w in weights]
    keep = []
    for l, planes in enumerate(grids):
        if len(planes) == 3:          # no time planes: the reference regularises nothing (gaussian_model.py:544-545)
            raise ValueError("3-plane (static) grids are not regularised by the reference")
        # res = [rx, ry, rz, rt]: plane k = (a, b) is [1, C, res[b], res[a]]
        a.res[l][0], a.res[l][1] = int(planes[0].shape[3]), int(planes[0].shape[2])
        a.res[l][2], a.res[l][3] = int(planes[1].shape[2]), int(planes[2].shape[2])
        for k, p in enumerate(planes):
            t = p.detach()
            if not t.is_cuda or t.dtype != torch.float32:
                raise RuntimeError("fdgs regularisers need CUDA float32 planes (there is no CPU fallback)")
            if not t.is_contiguous():
                t = t.contiguous()
            keep.append(t)
            a.planes[l][k] = t.data_ptr()
            a.d_planes[l][k] = grad_ptrs[l][k]
    return keep


def _launch(grids, weights, grad_ptrs, loss, grad_scale=None):
    lib = L.load()
    a = L.FdgsPlaneReg()
    keep = _fill(a, grids, weights, grad_ptrs)
    a.grad_scale = grad_scale.data_ptr() if grad_scale is not None else None
    a.loss = loss.data_ptr() if loss is not None else None
    dev = grids[0][0].device
    with L.on_device(dev):
        L.check(lib.fdgs_plane_regulation(C.byref(a), L.stream_ptr()),
                "fdgs_plane_regulation")
    del keep


class _Regulation(torch.autograd.Function):
    @staticmethod
    def forward(ctx, weights, levels, *planes):
        grids = [list(planes[6 * l:6 * l + 6]) for l in range(levels)]
        dev = planes[0].device
        loss = torch.zeros((), dtype=torch.float32, device=dev)
        need = any(ctx.needs_input_grad[2:])
        grads = [[torch.zeros_like(p, memory_format=torch.contiguous_format) if need else None for p in lv] for lv in grids]
        ptrs = [[g.data_ptr() if g is not None else None for g in lv] for lv in grads]
        _launch(grids, weights, ptrs, loss)
        ctx.grads = [g for lv in grads for g in lv]
        return loss

    @staticmethod
    def backward(ctx, g):
        return (None, None) + tuple(None if d is None else d * g for d in ctx.grads)


def compute_regulation(grids, time_smoothness_weight, l1_time_planes_weight, plane_tv_weight):
    """gaussian_model.py:576-577 (same argument order).  `grids`: deformation_net.grid.grids (or the network)."""
    gl = _grids_of(grids)
    flat = [p for lv in gl for p in lv]
    return _Regulation.apply((plane_tv_weight, time_smoothness_weight, l1_time_planes_weight), len(gl), *flat)


def plane_regulation(grids):
    """gaussian_model.py:538-549 _plane_regulation: sum of compute_plane_smoothness over planes 0, 1, 3 of every level."""
    return compute_regulation(grids, 0.0, 0.0, 1.0)


def time_regulation(grids):
    """gaussian_model.py:550-561 _time_regulation: the same over the space-time planes 2, 4, 5."""
    return compute_regulation(grids, 1.0, 0.0, 0.0)


def l1_regulation(grids):
    """gaussian_model.py:562-575 _l1_regulation: sum of mean |1 - plane| over the space-time planes."""
    return compute_regulation(grids, 0.0, 1.0, 0.0)


class _PlaneSmoothness(torch.autograd.Function):
    @staticmethod
    def forward(ctx, t):
        lib = L.load()
        if t.dim() != 4 or not t.is_cuda or t.dtype != torch.float32:
            raise RuntimeError("compute_plane_smoothness expects a CUDA float32 [B, C, H, W] tensor")
        tc = t.contiguous()
        B, Cc, H, W = tc.shape
        loss = torch.zeros((), dtype=torch.float32, device=t.device)
        grad = torch.zeros_like(tc) if ctx.needs_input_grad[0] else None
        with L.on_device(t.device):
            L.check(lib.fdgs_plane_smoothness(tc.data_ptr(), grad.data_ptr() if grad is not None else None, int(B * Cc), int(H),
                                              int(W), 1.0, None, loss.data_ptr(),
                                              L.stream_ptr()), "fdgs_plane_smoothness")
        ctx.grad = grad
        return loss

    @staticmethod
    def backward(ctx, g):
        return ctx.grad * g


def compute_plane_smoothness(t):
    """scene/regulation.py:22-28: mean of the squared second difference along dim 2 of a [B, C, H, W] plane."""
    return _PlaneSmoothness.apply(t)


def add_regulation_(grids, time_smoothness_weight, l1_time_planes_weight, plane_tv_weight, loss=None):
    """Fused training path: returns the regulariser value (a 0-d tensor; accumulated into `loss` when given) and ADDS its
    gradient to the planes' existing .grad tensors in the same kernel launch (weight 1 in the total loss, train.py:251-254).
    Every plane must already own a contiguous float32 .grad (e.g. a fdgs.dp.GradBucket slot)."""
    gl = _grids_of(grids)
    ptrs = []
    for lv in gl:
        row = []
        for p in lv:
            if p.grad is None or not p.grad.is_contiguous() or p.grad.dtype != torch.float32 or p.grad.shape != p.shape:
                raise RuntimeError("add_regulation_ needs every plane to own a contiguous float32 .grad (see fdgs.dp.GradBucket)")
            row.append(p.grad.data_ptr())
        ptrs.append(row)
    if loss is None:
        loss = torch.zeros((), dtype=torch.float32, device=gl[0][0].device)
    _launch(gl, (plane_tv_weight, time_smoothness_weight, l1_time_planes_weight), ptrs, loss)
    return loss
