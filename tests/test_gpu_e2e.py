"""Whole hot path, end to end: one training iteration (deformation -> activations -> rasterize -> L1 -> backward over
all views, gradient bucket) through fdgs vs. the reference arm (the reference's PyTorch deformation ops,
oracle/ref_port.py, + the UNMODIFIED reference CUDA rasterizer built into oracle/_ref) on the same seeded scene.
This is the drop-in claim of INTEGRATION.md checked in one piece: same images, same loss, same gradients for every
Gaussian parameter, HexPlane plane and MLP weight."""
import pytest
import torch

import util

pytestmark = pytest.mark.gpu


def _run(impl, cfg_name, views, points, fused_loss):
    import bench
    import fdgs  # noqa: F401
    from fdgs.dp import train_iteration
    dev = "cuda:0"
    cfg, args, pc, net, cams, gts_host = bench.build_scene(cfg_name, impl, dev, views, points)
    bucket = bench.make_bucket(pc, net, dev)
    bg = torch.ones(3, device=dev)
    gts = [g.to(dev) for g in gts_host]
    if impl == "fdgs":
        from fdgs.render import render as render_fn
        loss_fn = None
        if fused_loss:
            from fdgs.losses import l1_view_loss as loss_fn
    else:
        import build_ref
        import ref_port
        from fdgs.diff_gaussian_rasterization import GaussianRasterizationSettings as SettingsTuple
        refC = build_ref.load()

        def render_fn(cam, pc_, bg_color=None, stage="fine"):
            return ref_port.ref_render(refC, SettingsTuple, cam, pc_, bg_color, stage=stage)
        loss_fn = None
    with torch.no_grad():
        images = [render_fn(c, pc, bg_color=bg, stage="fine")["render"].clone() for c in cams]
    loss, radii, _vis = train_iteration(render_fn, pc, cams, gts, bg, bucket, 0, 1, read_loss=True, loss_fn=loss_fn)
    grads = {"viewspace": bucket.extra["viewspace"].clone()}
    for name in ("_xyz", "_scaling", "_rotation", "_opacity", "_features_dc", "_features_rest"):
        grads[name] = getattr(pc, name).grad.clone()
    for k, p in net.named_parameters():
        if p.grad is not None:
            grads["net." + k] = p.grad.clone()
    return images, loss, radii.clone(), grads


@pytest.mark.parametrize("cfg_name,views,points", [("tiny", 2, 0), ("dynerf_small", 2, 0), ("dnerf50k", 2, 20000)])
def test_train_iteration_matches_reference_arm(cfg_name, views, points):
    import build_ref
    if not build_ref.available():
        pytest.skip("oracle/_ref (reference rasterizer built from /root/reference) is not in this snapshot")
    img_r, loss_r, radii_r, g_r = _run("reference", cfg_name, views, points, False)
    img_f, loss_f, radii_f, g_f = _run("fdgs", cfg_name, views, points, True)
    assert torch.equal(radii_f, radii_r)
    for a, b in zip(img_f, img_r):
        # (observed, profiles/r02_parity_e2e.log: 0 / 72 000 values at `tiny`, 3 / 72 000 at `dynerf_small`, 39 and 73 / 1 920 000 at
        #  `dnerf50k`, worst 2.8e-3 - the allowance numel / 20 000 is 1.3x the largest count seen)
        # images: 1e-5 bar + 1 ulp of the fused activations.  A Gaussian whose alpha sits within an ulp of the 1/255 skip
        # threshold may be taken on one side and skipped on the other (libm-level differences upstream): a handful of
        # pixels may then differ by up to 1/255 of a colour.
        d = (a - b).abs() / b.abs().max()
        util.plog("e2e", f"{cfg_name}: image {tuple(a.shape)}: pixels-values off by > 2e-5: {int((d > 2e-5).sum())} of {a.numel()} "
                         f"(allowance {max(4, a.numel() // 20000)}), worst {float(d.max()):.2e}")
        assert int((d > 2e-5).sum()) <= max(4, a.numel() // 20000), float(d.max())
        assert float(d.max()) < 5e-3
    assert abs(loss_f - loss_r) <= 2e-6 * abs(loss_r)
    assert set(g_f) == set(g_r)
    for k in sorted(g_r):
        # float-atomic sums on both sides (the reference's own run-to-run spread is ~1e-5 of the largest entry); the
        # few Gaussians whose alpha flips across the 1/255 threshold (see above) carry a different gradient on the two
        # sides, so the bar is: all but a handful of entries within 2e-4 of the largest, and 2e-3 in the L2 sense
        a, b = g_f[k].double(), g_r[k].double()
        d = (a - b).abs() / (b.abs().max() + 1e-30)
        util.plog("e2e", f"{cfg_name}: grad {k}: entries off by > 2e-4: {int((d > 2e-4).sum())} of {a.numel()}, worst {float(d.max()):.2e}, "
                         f"L2 {float((a - b).norm() / (b.norm() + 1e-30)):.2e}")
        assert int((d > 2e-4).sum()) <= max(8, a.numel() // 2000), (k, float(d.max()))
        assert float((a - b).norm() / (b.norm() + 1e-30)) < 2e-3, k


def test_view_stager_and_shared_features_change_nothing():
    """fdgs.dp.ViewStager (GT images uploaded on a copy stream, one event per view) and the once-per-iteration SH cat are
    scheduling changes only: same loss and SH gradients as device-resident targets / a cat per view (up to the run-to-run
    spread of the float atomics in the loss reduction and the blend backward)."""
    import bench
    import fdgs  # noqa: F401
    from fdgs.dp import train_iteration, ViewStager
    from fdgs.render import render as render_fn
    from fdgs.losses import l1_view_loss as loss_fn
    dev = "cuda:0"
    cfg, args, pc, net, cams, gts_host = bench.build_scene("tiny", "fdgs", dev, 3, 0)
    bucket = bench.make_bucket(pc, net, dev)
    bg = torch.ones(3, device=dev)
    gts = [g.to(dev) for g in gts_host]
    res = []
    for targets, share in ((gts, False), (gts, True), ("stager", True)):
        if targets == "stager":
            targets = ViewStager(gts_host, dev).upload(gts_host, range(len(gts_host)))
            assert targets.nbytes == gts_host[0].numel() * 4           # the first view's copy is queued by upload() ...
            for i, g in enumerate(gts):
                assert torch.equal(targets[i], g)
            assert targets.nbytes == sum(g.numel() * 4 for g in gts_host)   # ... the others as their predecessor is fetched
            targets.upload(gts_host, range(len(gts_host)))
            assert torch.equal(targets[2], gts[2]) and torch.equal(targets[0], gts[0])     # out-of-order access
            targets.upload(gts_host, range(len(gts_host)))
        loss, _, _ = train_iteration(render_fn, pc, cams, targets, bg, bucket, 0, 1, read_loss=True, loss_fn=loss_fn,
                                  share_features=share)
        res.append((loss, pc._features_dc.grad.clone(), pc._features_rest.grad.clone()))
    assert float(res[0][1].abs().sum()) > 0 and float(res[0][2].abs().sum()) > 0
    for r in res[1:]:
        assert abs(r[0] - res[0][0]) <= 1e-6 * abs(res[0][0])
        for a, b in ((r[1], res[0][1]), (r[2], res[0][2])):
            assert float((a - b).abs().max()) <= 1e-4 * float(b.abs().max())


def test_speculative_iteration_equals_exact_and_redoes_on_overflow():
    """train_iteration(speculative=True): same loss / gradients as the exact path, and a step whose instance workspace
    was too small is redone before its gradients are used."""
    import bench
    import fdgs  # noqa: F401
    from fdgs import diff_gaussian_rasterization as R
    from fdgs.dp import train_iteration
    from fdgs.render import render as render_fn
    from fdgs.losses import l1_view_loss as loss_fn
    dev = "cuda:0"
    cfg, args, pc, net, cams, gts_host = bench.build_scene("dynerf_small", "fdgs", dev, 3, 0)
    bucket = bench.make_bucket(pc, net, dev)
    bg = torch.ones(3, device=dev)
    gts = [g.to(dev) for g in gts_host]
    R.check_pending()
    l0, r0, v0 = train_iteration(render_fn, pc, cams, gts, bg, bucket, 0, 1, read_loss=True, loss_fn=loss_fn)
    f0 = bucket.flat.clone()
    l1, r1, v1 = train_iteration(render_fn, pc, cams, gts, bg, bucket, 0, 1, read_loss=True, loss_fn=loss_fn, speculative=True)
    assert abs(l1 - l0) <= 1e-6 * abs(l0) and torch.equal(r0, r1) and torch.equal(v0, v1)
    assert util.rel_err(bucket.flat, f0) < 1e-5
    for k in list(R._caps):                                           # force an overflow in the next speculative step
        R._caps[k] = max(R._caps[k] // 50, 1)
    R.SPEC_SLACK, slack = 0, R.SPEC_SLACK
    try:
        l2, r2, v2 = train_iteration(render_fn, pc, cams, gts, bg, bucket, 0, 1, read_loss=True, loss_fn=loss_fn, speculative=True)
    finally:
        R.SPEC_SLACK = slack
    assert abs(l2 - l0) <= 1e-6 * abs(l0) and torch.equal(r0, r2)
    assert util.rel_err(bucket.flat, f0) < 1e-5


@pytest.mark.parametrize("views", [1, 3])
def test_early_identity_gradients_equal_the_autograd_route(views):
    """train_iteration(overlap_sh=True): scale / rotation / opacity / SH gradients of the last view are added into their
    bucket slots by the deformation backward's identity sink (before its kernels run; where the early all-reduce starts at
    world > 1) instead of by autograd after it.  Same additions: the bucket is identical up to float-atomic noise."""
    import bench
    import fdgs  # noqa: F401
    from fdgs.dp import train_iteration
    from fdgs.render import render as render_fn
    from fdgs.losses import l1_view_loss as loss_fn
    dev = "cuda:0"
    cfg, args, pc, net, cams, gts_host = bench.build_scene("dynerf_small", "fdgs", dev, views, 0)
    bucket = bench.make_bucket(pc, net, dev)
    assert bucket.span([pc._scaling, pc._rotation, pc._opacity, pc._features_dc, pc._features_rest]) is not None
    bg = torch.ones(3, device=dev)
    gts = [g.to(dev) for g in gts_host]
    l0, _, _ = train_iteration(render_fn, pc, cams, gts, bg, bucket, 0, 1, read_loss=True, loss_fn=loss_fn)
    f0 = bucket.flat.clone()
    l1, _, _ = train_iteration(render_fn, pc, cams, gts, bg, bucket, 0, 1, read_loss=True, loss_fn=loss_fn, overlap_sh=True)
    assert net.deformation_net.identity_sink is None
    assert abs(l1 - l0) <= 1e-6 * abs(l0)
    assert util.rel_err(bucket.flat, f0) < 1e-5
    for p in (pc._scaling, pc._rotation, pc._opacity, pc._features_dc, pc._features_rest):
        assert float(p.grad.abs().sum()) > 0
