"""TEST INFRASTRUCTURE — CPU restatements (numpy) of the reference's init-time native pieces.

  knn_mean_dist2      submodules/simple-knn/simple_knn.cu:134-192: for every point the three smallest
                      d.x*d.x + d.y*d.y + d.z*d.z over the OTHER points (FP32, fused like nvcc contracts it), mean = sum / 3.
                      The reference's Morton boxes only prune the search; the result is the exact 3-NN, which brute force gives.
  voxel_downsample    utils/grid_pruning.py:16-41 -> open3d 0.17 PointCloud::VoxelDownSample (geometry/PointCloud.cpp:
                      voxel_min_bound = min_bound - voxel_size * 0.5; index = floor((p - voxel_min_bound) / voxel_size);
                      AccumulatedPoint averages positions and colours in double).  open3d is not installed in this image:
                      its published algorithm is restated.  Parity unpinned for this function (no reference run possible here).
Never imported by the product."""
import numpy as np


def knn_mean_dist2(points):
    p = np.asarray(points, dtype=np.float32)
    P = p.shape[0]
    out = np.empty(P, dtype=np.float32)
    for i0 in range(0, P, 512):
        blk = p[i0:i0 + 512]
        d = blk[:, None, :] - p[None, :, :]                      # FP32 differences (other - self differs only in sign)
        # fma(dz, dz, fma(dy, dy, dx * dx)) emulated in double then rounded at each step
        t = (d[..., 0].astype(np.float64) * d[..., 0]).astype(np.float32)
        t = (d[..., 1].astype(np.float64) * d[..., 1] + t).astype(np.float32)
        t = (d[..., 2].astype(np.float64) * d[..., 2] + t).astype(np.float32)
        t[np.arange(blk.shape[0]), i0 + np.arange(blk.shape[0])] = np.float32(np.finfo(np.float32).max)
        b = np.sort(np.partition(t, 2, axis=1)[:, :3], axis=1)
        out[i0:i0 + 512] = ((b[:, 0] + b[:, 1]) + b[:, 2]) / np.float32(3.0)
    return out


def voxel_downsample(points, colors, voxel_size):
    p = np.asarray(points, dtype=np.float32).astype(np.float64)
    c = None if colors is None else np.asarray(colors, dtype=np.float32).astype(np.float64)
    vmin = p.min(axis=0) - voxel_size * 0.5
    idx = np.floor((p - vmin) * (1.0 / voxel_size)).astype(np.int64)
    uniq, inv, cnt = np.unique(idx, axis=0, return_inverse=True, return_counts=True)
    inv = inv.reshape(-1)
    sp = np.zeros((uniq.shape[0], 3)); np.add.at(sp, inv, p)
    outp = (sp / cnt[:, None]).astype(np.float32)
    outc = None
    if c is not None:
        sc = np.zeros((uniq.shape[0], 3)); np.add.at(sc, inv, c)
        outc = (sc / cnt[:, None]).astype(np.float32)
    order = np.lexsort((outp[:, 2], outp[:, 1], outp[:, 0]))
    return outp[order], (None if outc is None else outc[order])
