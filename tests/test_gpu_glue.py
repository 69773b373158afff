"""Fused activation glue (include/fdgs.h: fdgs_activate_*) against the ATen ops the reference calls
(gaussian_renderer/__init__.py:97-99): torch.exp, F.normalize(dim=1, eps=1e-12), torch.sigmoid."""
import pytest
import torch

import util

pytestmark = pytest.mark.gpu


def _ref(s, r, o):
    return torch.exp(s), torch.nn.functional.normalize(r), torch.sigmoid(o)


@pytest.mark.parametrize("P", [1, 257, 100_003])
def test_fused_activations_match_aten(P):
    from fdgs.activations import fused_activations
    gen = torch.Generator(device="cuda").manual_seed(P)
    s = (torch.randn(P, 3, generator=gen, device="cuda") - 4).requires_grad_(True)
    r = torch.randn(P, 4, generator=gen, device="cuda").requires_grad_(True)
    o = (2 * torch.randn(P, 1, generator=gen, device="cuda")).requires_grad_(True)
    outs = fused_activations(s, r, o)
    s2, r2, o2 = (t.detach().clone().requires_grad_(True) for t in (s, r, o))
    refs = _ref(s2, r2, o2)
    for a, b in zip(outs, refs):
        assert a.shape == b.shape
        torch.testing.assert_close(a, b, rtol=2e-6, atol=1e-7)       # 1-2 ulp: expf / division order
    gs = [torch.randn(t.shape, generator=gen, device="cuda") for t in outs]
    torch.autograd.backward(outs, gs)
    torch.autograd.backward(refs, gs)
    for a, b in ((s, s2), (r, r2), (o, o2)):
        torch.testing.assert_close(a.grad, b.grad, rtol=1e-5, atol=1e-6)


def test_fused_activations_degenerate_rows_and_partial_grads():
    from fdgs.activations import fused_activations
    s = torch.zeros(4, 3, device="cuda", requires_grad=True)
    r = torch.tensor([[0., 0, 0, 0], [1e-20, 0, 0, 0], [0, 3, 0, 4], [1, 1, 1, 1]], device="cuda", requires_grad=True)
    o = torch.tensor([[-100.], [0.], [100.], [1.]], device="cuda")         # no grad wanted for opacity
    so, ro, oo = fused_activations(s, r, o)
    rs, rr, ro_ref = _ref(s.detach(), r.detach(), o)
    torch.testing.assert_close(ro, rr, rtol=2e-6, atol=0)
    torch.testing.assert_close(oo, ro_ref, rtol=2e-6, atol=1e-38)
    assert not oo.requires_grad
    (ro.sum() + so.sum()).backward()
    r2 = r.detach().clone().requires_grad_(True)
    torch.nn.functional.normalize(r2).sum().backward()
    torch.testing.assert_close(r.grad, r2.grad, rtol=1e-5, atol=1e-6)      # includes the clamp_min(eps) rows
    assert torch.equal(s.grad, torch.ones_like(s))


def test_render_uses_fused_activations_and_matches_aten_path():
    """render() with the reference's activations (fused) vs. the same model with wrapped callables (ATen path)."""
    import fdgs  # noqa: F401
    from fdgs import synth
    from fdgs.configs import CONFIGS
    from fdgs.deformation import deform_network
    from fdgs.render import render
    cfg = CONFIGS["tiny"]
    net = synth.seeded_deformation(deform_network, cfg["args"](), seed=5, device="cuda")
    g = synth.make_gaussians(cfg["P"], extent=cfg["extent"], logscale=cfg["logscale"], seed=1, device="cuda")
    pc = synth.SynthGaussians(g, net).to("cuda")
    net.deformation_net.set_aabb(pc._xyz.detach().max(0).values.tolist(), pc._xyz.detach().min(0).values.tolist())
    cam = synth.make_cameras(cfg["W"], cfg["H"], cfg["fovx"], 1, device="cuda")[0]
    bg = torch.ones(3, device="cuda")
    out_fused = render(cam, pc, bg_color=bg)["render"]
    out_fused.sum().backward()
    g_fused = pc._scaling.grad.clone(), pc._rotation.grad.clone(), pc._opacity.grad.clone()
    for p in (pc._scaling, pc._rotation, pc._opacity, pc._xyz, pc._features_dc, pc._features_rest):
        p.grad = None
    net.zero_grad(set_to_none=True)
    pc.scaling_activation = lambda t: torch.exp(t)           # not the reference's function objects -> ATen path
    out_aten = render(cam, pc, bg_color=bg)["render"]
    out_aten.sum().backward()
    torch.testing.assert_close(out_fused, out_aten, rtol=1e-5, atol=1e-6)
    # the rasterizer backward sums with float atomics: two runs of the SAME path differ by ~1e-5 of the largest gradient
    for a, b in zip(g_fused, (pc._scaling.grad, pc._rotation.grad, pc._opacity.grad)):
        assert util.rel_err(a, b) < 1e-4


@pytest.mark.parametrize("shape", [(3, 64, 200), (3, 1014, 1352), (7,), (1, 5, 3)])
def test_fused_l1_loss_matches_aten(shape):
    """fdgs.losses.l1_loss vs utils/loss_utils.py:20-21 (torch.abs(a - b).mean()), value and both gradients."""
    from fdgs.losses import l1_loss, l1_view_loss
    gen = torch.Generator(device="cuda").manual_seed(len(shape))
    a = torch.rand(shape, generator=gen, device="cuda").requires_grad_(True)
    b = torch.rand(shape, generator=gen, device="cuda").requires_grad_(True)
    with torch.no_grad():
        b[..., :1] = a[..., :1]                                   # exact ties: sign(0) = 0
    a2, b2 = a.detach().clone().requires_grad_(True), b.detach().clone().requires_grad_(True)
    l = l1_loss(a, b)
    l_ref = torch.abs(a2 - b2).mean()
    torch.testing.assert_close(l, l_ref, rtol=1e-5, atol=0)
    (3.0 * l).backward()
    (3.0 * l_ref).backward()
    torch.testing.assert_close(a.grad, a2.grad, rtol=1e-6, atol=0)
    torch.testing.assert_close(b.grad, b2.grad, rtol=1e-6, atol=0)
    lv = l1_view_loss(a.detach(), b.detach(), 8)
    torch.testing.assert_close(lv, l_ref.detach() / 8, rtol=1e-5, atol=0)


def test_activations_folded_into_the_rasterizer_equal_the_separate_kernels():
    """rasterize_gaussians_raw (exp / normalize / sigmoid and their VJPs inside k_preprocess / k_preprocess_bwd) vs
    fused_activations followed by the ordinary rasterizer call: the same device functions, so images, radii and depth are
    bit-identical and the raw-input gradients agree up to the blend backward's float atomics; and both agree with ATen."""
    import math
    import util
    from fdgs import synth
    from fdgs.activations import fused_activations
    from fdgs.diff_gaussian_rasterization import GaussianRasterizer, rasterize_gaussians_raw
    P, W, H = 20_000, 320, 200
    g = synth.make_gaussians(P, extent=1.3, logscale=-3.8, seed=4, device="cuda")
    g["rotation"][:5] = 0.0                                 # zero quaternions: the clamp_min(eps) branch of normalize
    cam = synth.make_camera(W, H, 0.69, angle=0.2, device="cuda")
    sc = dict(W=W, H=H, tanfovx=math.tan(cam.FoVx * 0.5), tanfovy=math.tan(cam.FoVy * 0.5), bg=torch.ones(3, device="cuda"),
              viewmatrix=cam.world_view_transform, projmatrix=cam.full_proj_transform, campos=cam.camera_center, sh_degree=3)
    settings = util.settings(sc)
    shs0 = torch.cat([g["f_dc"], g["f_rest"]], dim=1)
    gimg = torch.randn((3, H, W), generator=torch.Generator().manual_seed(2)).cuda()

    def run(mode):
        leaves = {k: g[k].clone().requires_grad_(True) for k in ("xyz", "scaling", "rotation", "opacity")}
        shs = shs0.clone().requires_grad_(True)
        m2d = torch.zeros((P, 3), device="cuda", requires_grad=True)
        if mode == "folded":
            color, radii, depth = rasterize_gaussians_raw(leaves["xyz"], m2d, shs, leaves["opacity"], leaves["scaling"],
                                                          leaves["rotation"], settings)
        else:
            if mode == "fused":
                s, r, o = fused_activations(leaves["scaling"], leaves["rotation"], leaves["opacity"])
            else:
                s, r, o = torch.exp(leaves["scaling"]), torch.nn.functional.normalize(leaves["rotation"]), torch.sigmoid(leaves["opacity"])
            color, radii, depth = GaussianRasterizer(settings)(means3D=leaves["xyz"], means2D=m2d, opacities=o, shs=shs, scales=s,
                                                               rotations=r)
        color.backward(gimg)
        grads = {k: v.grad for k, v in leaves.items()}
        grads["shs"], grads["m2d"] = shs.grad, m2d.grad
        return color.detach(), radii, depth.detach(), grads

    cf, rf, df, gf = run("folded")
    cu, ru, du, gu = run("fused")
    ca, ra, da, ga = run("aten")
    assert torch.equal(cf, cu) and torch.equal(rf, ru) and torch.equal(df, du)
    assert int((rf != ra).sum()) <= 2 and util.rel_err(cf, ca) < 1e-4      # ATen's exp / sigmoid differ by an ulp
    for k in gf:
        assert util.rel_err(gf[k], gu[k]) < 2e-6, k
        assert util.rel_err(gf[k], ga[k]) < 2e-4, k
