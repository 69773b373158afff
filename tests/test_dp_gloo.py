"""Host logic of the view-parallel path (fdgs/dp.py) on CPU: world_size-2 `gloo` processes.

The render function is a small differentiable stand-in (no CUDA): what is under test is the sharding rule
(view i -> rank i mod G), the loss scaling that makes summed shard gradients equal the reference's single
l1_loss over the concatenated batch (train.py:231-244), the flat bucket whose slots ARE the .grad tensors,
and the single all-reduce."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import fdgs  # noqa: F401
from fdgs import dp


class _PC(torch.nn.Module):
    def __init__(self):
        super().__init__()
        g = torch.Generator().manual_seed(3)
        self.a = torch.nn.Parameter(torch.randn(37, 3, generator=g))
        self.b = torch.nn.Parameter(torch.randn(5, generator=g))
        self.frozen = torch.nn.Parameter(torch.randn(4, generator=g), requires_grad=False)


def _render(cam, pc, bg_color=None, stage="fine"):
    vs = torch.zeros(37, 3, requires_grad=True)
    img = (pc.a * cam).sum(0)[:, None, None] * torch.ones(3, 4, 6) + pc.b.sum() * cam + bg_color[:, None, None] + vs.sum() * cam
    vs.retain_grad()
    # radii differ per view AND per Gaussian; Gaussian 36 is invisible in every view, Gaussian k only in view k
    radii = torch.full((37,), int(cam * 10), dtype=torch.int32)
    radii[36] = 0
    k = int(round((cam - 0.3) / 0.2))
    radii[:8] = 0
    if k < 8:
        radii[k] = 50 + k
    return dict(render=img, viewspace_points=vs, radii=radii)


def _batch(B):
    g = torch.Generator().manual_seed(11)
    cams = [0.3 + 0.2 * i for i in range(B)]
    gts = [torch.rand(3, 4, 6, generator=g) for _ in range(B)]
    return cams, gts


def _single_process_reference(B):
    pc = _PC()
    cams, gts = _batch(B)
    bg = torch.ones(3)
    outs = [_render(c, pc, bg)["render"] for c in cams]
    loss = (torch.cat(outs, 0) - torch.cat(gts, 0)).abs().mean()       # the reference's l1_loss(cat, cat)
    loss.backward()
    return float(loss), pc.a.grad.clone(), pc.b.grad.clone()


def _worker(rank, world, port, B, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        pc = _PC()
        bucket = dp.GradBucket(pc.parameters(), extra={"viewspace": (37, 3)}, device="cpu")
        assert pc.a.grad.data_ptr() == bucket.views[0].data_ptr()
        cams, gts = _batch(B)
        loss, radii, vis = dp.train_iteration(_render, pc, cams, gts, torch.ones(3), bucket, rank=rank, world=world)
        tot = loss.clone()
        dist.all_reduce(tot)
        # numpy copies, not tensors: a tensor travels through the queue as a handle served by THIS process, which may have
        # exited by the time the parent opens it (FileNotFoundError in multiprocessing.connection, seen once in ~10 runs)
        q.put((rank, float(tot), pc.a.grad.numpy().copy(), pc.b.grad.numpy().copy(), bucket.extra["viewspace"].numpy().copy(),
               radii.numpy().copy(), bucket.nbytes(), vis.numpy().copy()))
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_shard_views_rule():
    assert dp.shard_views(8, 0, 1) == list(range(8))
    assert dp.shard_views(8, 1, 2) == [1, 3, 5, 7]
    assert dp.shard_views(8, 3, 4) == [3, 7]
    assert dp.shard_views(8, 7, 8) == [7]
    assert dp.shard_views(2, 3, 4) == []          # more ranks than views: the rank idles but still joins the reduce
    assert sorted(sum((dp.shard_views(5, r, 3) for r in range(3)), [])) == list(range(5))


def test_bucket_slots_are_the_grads_and_skip_frozen():
    pc = _PC()
    b = dp.GradBucket(pc.parameters(), extra={"viewspace": (37, 3)}, device="cpu")
    assert len(b.params) == 2 and pc.frozen.grad is None
    assert all(o % 64 == 0 for o in b.offsets)
    (pc.a.sum() * 2 + pc.b.sum()).backward()
    assert float(b.flat.sum()) == pytest.approx(2 * 37 * 3 + 5)
    pc.a.grad = None                               # what optimizer.zero_grad(set_to_none=True) does
    b.zero()
    assert pc.a.grad.data_ptr() == b.views[0].data_ptr() and float(b.flat.abs().sum()) == 0.0
    assert b.all_reduce() is None                  # no process group: no-op


@pytest.mark.parametrize("B", [4, 3])
def test_world2_gloo_matches_single_process(B):
    ref_loss, ga, gb = _single_process_reference(B)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, B, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    res = [tuple(torch.from_numpy(x) if hasattr(x, "dtype") and not isinstance(x, torch.Tensor) else x for x in r) for r in res]
    cams, _ = _batch(B)
    per_view = torch.stack([_render(c, _PC(), torch.ones(3))["radii"] for c in cams])
    for rank, loss, a, b, vs, radii, nbytes, vis in res:
        assert loss == pytest.approx(ref_loss, rel=1e-6)
        assert torch.allclose(a, ga, rtol=1e-5, atol=1e-7) and torch.allclose(b, gb, rtol=1e-5, atol=1e-7)
        assert nbytes == 4 * (128 + 64 + 128)      # 111 -> 128, 5 -> 64, 111 -> 128 floats
        # densification inputs are those of the WHOLE batch on every rank (train.py:229-230): max / any over all B views
        assert torch.equal(radii, per_view.max(0).values)
        assert torch.equal(vis, (per_view > 0).any(0))
    assert not bool(res[0][7][36]) and bool(res[0][7][:B].all()) and not bool(res[0][7][B:8].any())
    assert torch.equal(res[0][4], res[1][4])       # the view-space accumulator is reduced too
    assert float(res[0][4].abs().sum()) > 0


class _PCFeat(_PC):
    """_PC plus the reference's split SH storage: get_features cats on every access (scene/gaussian_model.py:123-127)."""

    def __init__(self):
        super().__init__()
        g = torch.Generator().manual_seed(5)
        self._features_dc = torch.nn.Parameter(torch.randn(37, 1, 3, generator=g))
        self._features_rest = torch.nn.Parameter(torch.randn(37, 15, 3, generator=g))
        self.cats = 0

    @property
    def get_features(self):
        self.cats += 1
        return torch.cat((self._features_dc, self._features_rest), dim=1)


def _render_feat(cam, pc, bg_color=None, stage="fine"):
    out = _render(cam, pc, bg_color, stage)
    w = torch.linspace(0.1, 1.0, 16)[None, :, None] * cam
    out["render"] = out["render"] + (pc.get_features * w).sum((0, 1))[:, None, None]
    return out


@pytest.mark.parametrize("B", [1, 4])
def test_shared_features_gradients_equal_per_view_cat(B):
    cams, gts = _batch(B)
    res = []
    for share in (False, True):
        pc = _PCFeat()
        bucket = dp.GradBucket(pc.parameters(), extra={"viewspace": (37, 3)}, device="cpu")
        loss, _, _ = dp.train_iteration(_render_feat, pc, cams, gts, torch.ones(3), bucket, share_features=share)
        slot = {id(p): v for p, v in zip(bucket.params, bucket.views)}
        assert pc._features_dc.grad.data_ptr() == slot[id(pc._features_dc)].data_ptr()
        assert pc._features_rest.grad.data_ptr() == slot[id(pc._features_rest)].data_ptr()
        res.append((float(loss), pc.cats, [p.grad.clone() for p in bucket.params]))
    assert res[0][1] == B and res[1][1] == 1          # one cat per view vs one per iteration
    assert res[0][0] == res[1][0]
    for g0, g1 in zip(res[0][2], res[1][2]):
        assert torch.equal(g0, g1)                    # same additions in the same order: bit-identical
    assert pc._features_rest.grad.abs().sum() > 0


def _worker_feat(rank, world, port, B, overlap, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        pc = _PCFeat()
        bucket = dp.GradBucket(pc.parameters(), extra={"viewspace": (37, 3)}, device="cpu")
        assert bucket.span([pc._features_dc, pc._features_rest]) is not None
        assert bucket.span([pc._features_rest, pc._features_dc]) is None
        calls = []
        orig = bucket.all_reduce_range
        bucket.all_reduce_range = lambda lo, hi, **kw: (calls.append((lo, hi, kw.get("async_op", False))), orig(lo, hi, **kw))[1]
        cams, gts = _batch(B)
        dp.train_iteration(_render_feat, pc, cams, gts, torch.ones(3), bucket, rank=rank, world=world, overlap_sh=overlap)
        q.put((rank, bucket.flat.numpy().copy(), calls))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("B", [4, 2, 1])         # 2 views per rank: shared SH leaf; 1 view per rank: per-parameter hooks;
def test_world2_early_sh_all_reduce_equals_single_collective(B):     # B = 1 < world: rank 1 owns no view, same collectives
    ctx = mp.get_context("spawn")
    flats = {}
    for overlap in (False, True):
        q = ctx.Queue()
        port = _free_port()
        procs = [ctx.Process(target=_worker_feat, args=(r, 2, port, B, overlap, q)) for r in range(2)]
        for p in procs:
            p.start()
        res = [q.get(timeout=120) for _ in procs]
        for p in procs:
            p.join(timeout=60)
            assert p.exitcode == 0
        res = [(r[0], torch.from_numpy(r[1]), r[2]) for r in res]
        assert torch.equal(res[0][1], res[1][1])
        flats[overlap] = res[0][1]
        if overlap:
            for r in res:                                         # EVERY rank, with or without views, issues the same schedule
                calls = r[2]
                assert len(calls) == 3 and calls[0][2] is True    # SH span first and asynchronous, then the two remainders
                assert calls[1][0] == 0 and calls[1][1] == calls[0][0] and calls[2][0] == calls[0][1]
            assert res[0][2] == res[1][2]
        else:
            assert res[0][2] == []
    assert torch.equal(flats[False], flats[True])                 # two-rank sums: order-independent, bit-identical
    assert float(flats[True].abs().sum()) > 0
