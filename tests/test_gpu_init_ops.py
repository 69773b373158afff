"""GPU parity of the init-time pieces against their CPU restatements (oracle/init_oracle.py):
distCUDA2 (exact 3-NN mean squared distance: bit-exact) and the Instant4D voxel down-sampling."""
import numpy as np
import pytest
import torch

import fdgs  # noqa: F401
from fdgs import init_ops
import init_oracle

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("P,kind", [(4, "uniform"), (1000, "uniform"), (20_011, "uniform"), (8000, "clustered"), (3000, "duplicates")])
def test_dist_cuda2_is_the_exact_3nn(P, kind):
    rng = np.random.default_rng(P)
    pts = rng.uniform(-1.3, 1.3, size=(P, 3)).astype(np.float32)
    if kind == "clustered":                      # most points in a tiny blob, a few far away: deep shell searches
        pts[: P - 50] = (rng.normal(0, 1e-3, size=(P - 50, 3)) + 0.7).astype(np.float32)
    if kind == "duplicates":                     # exact duplicates: zero distances
        pts[: P // 2] = pts[P // 2: 2 * (P // 2)]
    got = init_ops.distCUDA2(torch.tensor(pts).cuda()).cpu().numpy()
    want = init_oracle.knn_mean_dist2(pts)
    assert got.dtype == np.float32 and got.shape == (P,)
    assert np.array_equal(got, want), float(np.abs(got - want).max())


def test_dist_cuda2_errors_and_empty():
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        init_ops.distCUDA2(torch.zeros(5, 3))
    assert init_ops.distCUDA2(torch.zeros(0, 3, device="cuda")).shape == (0,)


@pytest.mark.parametrize("P,voxel", [(5000, 0.2), (50_000, 0.05), (1000, 10.0), (7, 1e-3)])
def test_voxel_downsample_matches_open3d_algorithm(P, voxel):
    rng = np.random.default_rng(P + 1)
    pts = rng.uniform(-2, 2, size=(P, 3)).astype(np.float32)
    cols = rng.uniform(0, 1, size=(P, 3)).astype(np.float32)
    gp, gc = init_ops.voxel_downsample(torch.tensor(pts).cuda(), torch.tensor(cols).cuda(), voxel)
    wp, wc = init_oracle.voxel_downsample(pts, cols, voxel)
    assert gp.shape[0] == wp.shape[0]
    # means accumulated in double on both sides, rounded once: equal up to the (order-dependent) last bit of the double sum
    np.testing.assert_allclose(gp.cpu().numpy(), wp, rtol=0, atol=1e-6)
    np.testing.assert_allclose(gc.cpu().numpy(), wc, rtol=0, atol=1e-6)
    gp2, gc2 = init_ops.voxel_downsample(torch.tensor(pts).cuda(), None, voxel)
    assert gc2 is None and torch.equal(gp2, gp)
