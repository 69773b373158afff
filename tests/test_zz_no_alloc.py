"""fdgs_knn_mean_dist2_static: the allocation-free variant of fdgs_knn_mean_dist2 (scratch in a file-scope device array of
libfdgs.so, include/fdgs.h).  Called straight through the C ABI and compared, bit for bit (the bar of tests/test_gpu_init_ops.py),
with the existing entry point on two calls in a row that share the array.  Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest
import torch

import fdgs  # noqa: F401
from fdgs import _lib as L

pytestmark = pytest.mark.gpu


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _cloud(P, seed, clustered=False):
    rng = np.random.default_rng(seed)
    pts = rng.uniform(-1.3, 1.3, size=(P, 3)).astype(np.float32)
    if clustered:
        pts[: P - 50] = (rng.normal(0, 1e-3, size=(P - 50, 3)) + 0.7).astype(np.float32)
    return torch.tensor(pts).cuda()


def _existing(lib, pts):
    P = pts.shape[0]
    out = torch.empty(P, dtype=torch.float32, device=pts.device)
    ws = torch.empty((lib.fdgs_knn_ws_bytes(P),), dtype=torch.uint8, device=pts.device)
    L.check(lib.fdgs_knn_mean_dist2(P, pts.data_ptr(), out.data_ptr(), ws.data_ptr(), ws.numel(), _stream()), "fdgs_knn_mean_dist2")
    torch.cuda.synchronize()
    return out


def test_static_scratch_knn_matches_the_workspace_entry_point_twice_in_a_row():
    lib = L.load()
    assert lib.fdgs_knn_static_max_points() == 262144
    # a large cloud first, then a smaller clustered one: the second call finds the first one's leftovers in the array
    clouds = [_cloud(150_001, 1), _cloud(8000, 2, clustered=True), _cloud(150_001, 1)]
    want = [_existing(lib, p) for p in clouds]
    outs = [torch.full((p.shape[0],), -1.0, dtype=torch.float32, device=p.device) for p in clouds]
    # set-up (once per device: the array's address is looked up) happens in this first small call, outside the measurement
    L.check(lib.fdgs_knn_mean_dist2_static(64, clouds[1].data_ptr(), outs[1].data_ptr(), _stream()), "fdgs_knn_mean_dist2_static")
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    torch_alloc0 = torch.cuda.memory_allocated()
    for p, o in zip(clouds, outs):              # back to back on one stream: calls that share the array must be serialised
        L.check(lib.fdgs_knn_mean_dist2_static(p.shape[0], p.data_ptr(), o.data_ptr(), _stream()), "fdgs_knn_mean_dist2_static")
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert torch.cuda.memory_allocated() == torch_alloc0
    assert free1 >= free0, (free0, free1)       # nothing was allocated on the device by the three calls
    for o, w in zip(outs, want):
        assert torch.equal(o, w), float((o - w).abs().max())
    assert torch.equal(outs[0], outs[2])


def test_static_scratch_knn_limits_and_errors():
    lib = L.load()
    P = lib.fdgs_knn_static_max_points()
    pts = _cloud(P, 3)
    out = torch.empty(P, dtype=torch.float32, device="cuda")
    L.check(lib.fdgs_knn_mean_dist2_static(P, pts.data_ptr(), out.data_ptr(), _stream()), "fdgs_knn_mean_dist2_static")
    torch.cuda.synchronize()
    assert torch.equal(out, _existing(lib, pts))                        # the stated maximum itself works
    rc = lib.fdgs_knn_mean_dist2_static(P + 1, pts.data_ptr(), out.data_ptr(), _stream())
    assert rc == -5                                                     # FDGS_E_TOO_LARGE, nothing launched
    assert b"exceeds the fixed maximum" in lib.fdgs_error_string(rc)
    with pytest.raises(RuntimeError, match="exceeds the fixed maximum"):
        L.check(rc, "fdgs_knn_mean_dist2_static")
    assert lib.fdgs_knn_mean_dist2_static(10, None, out.data_ptr(), _stream()) == -1      # FDGS_E_BADARG
    assert lib.fdgs_knn_mean_dist2_static(0, None, None, _stream()) == 0
    torch.cuda.synchronize()
