"""Init-time native pieces of the reference as CUDA kernels behind the C ABI.

  distCUDA2(points)                       drop-in for `from simple_knn._C import distCUDA2` (scene/gaussian_model.py:21,145):
                                          mean squared distance to the 3 nearest neighbours, float32 [P]
  voxel_downsample(points, colors, size)  utils/grid_pruning.py:16-41 (open3d `voxel_down_sample`): one averaged point and colour
                                          per occupied voxel; rows come back sorted by voxel index (open3d's order is that of an
                                          unordered_map, i.e. unspecified)

There is no CPU fallback: CPU tensors raise."""
import ctypes as C

import torch

from . import _lib as L


def _check(t, name, cols=3):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (there is no CPU fallback)")
    if t.dim() != 2 or t.shape[1] != cols:
        raise RuntimeError(f"{name} must have shape [P,{cols}]")
    return t.detach().to(torch.float32).contiguous()


def distCUDA2(points):
    lib = L.load()
    pts = _check(points, "points")
    P = pts.shape[0]
    out = torch.empty((P,), dtype=torch.float32, device=pts.device)
    if P == 0:
        return out
    ws = torch.empty((lib.fdgs_knn_ws_bytes(P),), dtype=torch.uint8, device=pts.device)
    with L.on_device(pts.device):
        L.check(lib.fdgs_knn_mean_dist2(P, pts.data_ptr(), out.data_ptr(), ws.data_ptr(), ws.numel(),
                                        L.stream_ptr()), "fdgs_knn_mean_dist2")
    return out


def voxel_downsample(points, colors, voxel_size):
    """Returns (points [V,3], colors [V,3] or None) as float32 CUDA tensors, rows in lexicographic (x, y, z) order."""
    lib = L.load()
    pts = _check(points, "points")
    cols = _check(colors, "colors") if colors is not None else None
    P = pts.shape[0]
    if cols is not None and cols.shape[0] != P:
        raise RuntimeError("points and colors must have the same number of rows")
    dev = pts.device
    out_p = torch.empty((P, 3), dtype=torch.float32, device=dev)
    out_c = torch.empty((P, 3), dtype=torch.float32, device=dev) if cols is not None else None
    n_out = torch.zeros((1,), dtype=torch.int32, device=dev)
    if P:
        ws = torch.empty((lib.fdgs_voxel_downsample_ws_bytes(P),), dtype=torch.uint8, device=dev)
        with L.on_device(dev):
            L.check(lib.fdgs_voxel_downsample(P, pts.data_ptr(), L.ptr(cols), float(voxel_size), out_p.data_ptr(), L.ptr(out_c), P,
                                              n_out.data_ptr(), ws.data_ptr(), ws.numel(),
                                              L.stream_ptr()), "fdgs_voxel_downsample")
    V = int(n_out.item())
    out_p = out_p[:V]
    order = torch.arange(V, device=dev)
    for a in (2, 1, 0):                      # stable lexicographic sort on (x, y, z)
        order = order[torch.sort(out_p[order, a], stable=True).indices]
    return out_p[order], (out_c[:V][order] if out_c is not None else None)
