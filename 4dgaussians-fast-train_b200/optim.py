"""Fused Adam and densification statistics of the training step.

  FusedAdam              drop-in for `torch.optim.Adam(l, lr=0.0, eps=1e-15)` as scene/gaussian_model.py:172-184 builds it
                         (param groups xyz / deformation / grid / f_dc / f_rest / opacity / scaling / rotation with their own
                         lr, betas (0.9, 0.999), no weight decay, no amsgrad).  Same `param_groups` / `state` / `state_dict()`
                         layout (state[p] = {"step", "exp_avg", "exp_avg_sq"}), so the reference's lr schedulers
                         (update_learning_rate, gaussian_model.py:211-230), its densification surgery on the optimizer state
                         (cat_tensors_to_optimizer / _prune_optimizer / replace_tensor_to_optimizer, :259-330) and checkpoints
                         work unchanged.  step() is ONE kernel launch over every tensor of every group (`fdgs_adam_step`);
                         its FP32 operation order is torch's foreach Adam's, so the results are bit-identical.
  add_densification_stats  train.py:346-349 + gaussian_model.py:521-523 in one launch (`fdgs_densify_stats`).

There is no CPU fallback: CPU parameters raise.
"""
import ctypes as C

import torch

from . import _lib as L


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, amsgrad=False):
        if weight_decay != 0 or amsgrad:
            raise NotImplementedError("FusedAdam implements the reference's configuration: weight_decay=0, amsgrad=False")
        if not 0.0 <= lr or not 0.0 <= eps or not 0.0 <= betas[0] < 1.0 or not 0.0 <= betas[1] < 1.0:
            raise ValueError("invalid Adam hyper-parameters")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=0, amsgrad=False))

    def _init_state(self, p):
        st = self.state[p]
        if len(st) == 0:
            st["step"] = torch.tensor(0.0, dtype=torch.float32)       # torch.optim.Adam keeps it as a CPU float32 scalar
            st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
            st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
        return st

    @torch.no_grad()
    def step(self, closure=None, zero_grad=False):
        """One Adam update of every parameter that has a gradient.  zero_grad=True also clears the gradients in the same
        pass (what `optimizer.zero_grad(set_to_none=False)` would do with one more sweep)."""
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        lib = L.load()
        by_key = {}
        for group in self.param_groups:
            b1, b2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                if not p.is_cuda:
                    raise RuntimeError("FusedAdam needs CUDA parameters (there is no CPU fallback)")
                if p.dtype != torch.float32 or p.grad.dtype != torch.float32:
                    raise RuntimeError("FusedAdam needs float32 parameters and gradients")
                if p.grad.is_sparse:
                    raise RuntimeError("FusedAdam does not support sparse gradients")
                if not (p.is_contiguous() and p.grad.is_contiguous()):
                    raise RuntimeError("FusedAdam needs contiguous parameters and gradients")
                st = self._init_state(p)
                if not (st["exp_avg"].is_contiguous() and st["exp_avg_sq"].is_contiguous()):
                    st["exp_avg"], st["exp_avg_sq"] = st["exp_avg"].contiguous(), st["exp_avg_sq"].contiguous()
                st["step"] += 1
                key = (p.device, float(b1), float(b2), float(group["eps"]))
                by_key.setdefault(key, []).append((p, st, float(group["lr"])))
        for (dev, b1, b2, eps), items in by_key.items():
            arr = (L.FdgsAdamTensor * len(items))()
            for t, (p, st, lr) in zip(arr, items):
                t.param, t.grad = p.data_ptr(), p.grad.data_ptr()
                t.exp_avg, t.exp_avg_sq = st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr()
                t.n, t.lr, t.step = p.numel(), lr, int(st["step"])
            a = L.FdgsAdamStep(ntensors=len(items), zero_grad=int(bool(zero_grad)), beta1=b1, beta2=b2, eps=eps, t=arr)
            with L.on_device(dev):
                L.check(lib.fdgs_adam_step(C.byref(a), L.stream_ptr()), "fdgs_adam_step")
        return loss


def add_densification_stats(radii, viewspace_grad, max_radii2D=None, xyz_gradient_accum=None, denom=None, visibility=None):
    """train.py:346-349 / gaussian_model.py:521-523 for one iteration, in place, visible Gaussians only:
        max_radii2D[v]        = max(max_radii2D[v], radii[v])
        xyz_gradient_accum[v] += norm(viewspace_grad[v, :2], dim=-1, keepdim=True)
        denom[v]              += 1
    radii: int32 [P] (max over the batch's views); viewspace_grad: [P,3] summed over views; visibility: bool [P] or None
    (= radii > 0, which is what train.py:230's any-over-views equals)."""
    lib = L.load()
    P = radii.shape[0]
    if not radii.is_cuda:
        raise RuntimeError("add_densification_stats needs CUDA tensors (there is no CPU fallback)")
    if radii.dtype != torch.int32 or not radii.is_contiguous():
        raise RuntimeError("radii must be a contiguous int32 tensor")
    vg = viewspace_grad
    if vg is not None and (vg.dtype != torch.float32 or not vg.is_contiguous() or vg.shape != (P, 3)):
        raise RuntimeError("viewspace_grad must be a contiguous float32 [P,3] tensor")
    for t in (max_radii2D, xyz_gradient_accum, denom):
        if t is not None and (t.dtype != torch.float32 or not t.is_contiguous() or t.numel() != P):
            raise RuntimeError("statistics tensors must be contiguous float32 with P elements")
    vis = None
    if visibility is not None:
        vis = visibility.contiguous().view(torch.uint8) if visibility.dtype == torch.bool else visibility.to(torch.uint8).contiguous()
    with L.on_device(radii.device):
        L.check(lib.fdgs_densify_stats(int(P), radii.data_ptr(), L.ptr(vis), L.ptr(vg), L.ptr(max_radii2D),
                                       L.ptr(xyz_gradient_accum), L.ptr(denom),
                                       L.stream_ptr()), "fdgs_densify_stats")
