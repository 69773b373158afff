"""GPU parity of the per-iteration training glue through the C ABI:
  * HexPlane regularisers (fdgs.regulation) vs the reference's own functions — scene/regulation.py
    compute_plane_smoothness loaded from the reference (baseline/_ref or /root/reference) when available, else the same
    tensor expressions restated here — value and gradient, 1e-6 relative;
  * FusedAdam vs torch.optim.Adam (the reference's optimizer, gaussian_model.py:184): parameters and both moments
    BIT-IDENTICAL over several steps with per-group lr, changing lr, a tensor without gradient, ragged sizes;
  * densification statistics vs the reference's tensor expressions (train.py:346-349, gaussian_model.py:521-523)."""
import pytest
import torch

import fdgs  # noqa: F401
from fdgs import optim as fopt
from fdgs import regulation as freg
from fdgs.configs import CONFIGS
from fdgs.deformation import deform_network
from fdgs import synth
import util

pytestmark = pytest.mark.gpu
VAL_TOL = 5e-6      # FP32 sums of up to 7e6 terms in two different orders (ATen's mean vs block partials kept in double)


def _ref_smoothness():
    try:
        import ref_deform
        return ref_deform.load_regulation().compute_plane_smoothness, "reference scene/regulation.py"
    except Exception:
        def smooth(t):      # scene/regulation.py:22-28 restated
            h = t.shape[2]
            first = t[..., 1:, :] - t[..., :h - 1, :]
            second = first[..., 1:, :] - first[..., :h - 2, :]
            return torch.square(second).mean()
        return smooth, "restated expression"


def _ref_compute_regulation(grids, tsw, l1w, ptw):
    """gaussian_model.py:538-577 over the reference's compute_plane_smoothness."""
    smooth, _ = _ref_smoothness()
    plane = sum(smooth(g[k]) for g in grids for k in (0, 1, 3))
    time = sum(smooth(g[k]) for g in grids for k in (2, 4, 5))
    l1 = sum(torch.abs(1 - g[k]).mean() for g in grids for k in (2, 4, 5))
    return ptw * plane + tsw * time + l1w * l1


@pytest.mark.parametrize("cfg", ["tiny", "dynerf", "hypernerf"])
def test_regulation_value_and_gradient_vs_reference(cfg):
    args = CONFIGS[cfg]["args"]()
    net = synth.seeded_deformation(deform_network, args, device="cuda", time_sigma=0.3)
    with torch.no_grad():       # the reference init leaves the spatial planes rough and the time planes near 1: perturb both
        g = torch.Generator(device="cuda").manual_seed(5)
        for lv in net.deformation_net.grid.grids:
            for p in lv:
                p.add_(torch.randn(p.shape, generator=g, device="cuda") * 0.1)
    grids = net.deformation_net.grid.grids
    w = (args.time_smoothness_weight, args.l1_time_planes, args.plane_tv_weight)
    ref = _ref_compute_regulation(grids, *w)
    ref.backward()
    gref = [p.grad.clone() for lv in grids for p in lv]
    for lv in grids:
        for p in lv:
            p.grad = None
    ours = freg.compute_regulation(grids, *w)
    (ours * 1.0).backward()
    assert abs(float(ours) - float(ref)) <= VAL_TOL * abs(float(ref)), (float(ours), float(ref))
    worst = 0.0
    for a, b in zip([p.grad for lv in grids for p in lv], gref):
        e = util.rel_err(a, b)
        worst = max(worst, e)
        assert e < 2e-6, e
    # the three partial sums and an upstream gradient other than 1
    for fn, wsel in ((freg.plane_regulation, (0.0, 0.0, 1.0)), (freg.time_regulation, (1.0, 0.0, 0.0)), (freg.l1_regulation, (0.0, 1.0, 0.0))):
        for lv in grids:
            for p in lv:
                p.grad = None
        r = _ref_compute_regulation(grids, *wsel)
        (r * 3.5).backward()
        gr = [None if p.grad is None else p.grad.clone() for lv in grids for p in lv]
        for lv in grids:
            for p in lv:
                p.grad = None
        o = fn(grids)
        (o * 3.5).backward()
        assert abs(float(o) - float(r)) <= VAL_TOL * abs(float(r))
        for p, b in zip([p for lv in grids for p in lv], gr):
            if b is None or not bool(b.abs().max() > 0):
                assert p.grad is None or not bool(p.grad.abs().max() > 0)
            else:
                assert util.rel_err(p.grad, b) < 2e-6
    # fused in-place path: value returned, gradient ADDED to existing .grad tensors (pre-filled at the gradient's own scale)
    for p, b in zip([p for lv in grids for p in lv], gref):
        p.grad = (0.5 * b).contiguous()
    v = freg.add_regulation_(grids, *w)
    assert abs(float(v) - float(ref)) <= VAL_TOL * abs(float(ref))
    for p, b in zip([p for lv in grids for p in lv], gref):
        assert util.rel_err(p.grad - 0.5 * b, b) < 5e-6
    util.plog("train_glue", f"regulation[{cfg}] vs {_ref_smoothness()[1]}: value {float(ours):.8e} vs {float(ref):.8e}, worst grad rel err {worst:.2e}")


def test_plane_smoothness_single_plane_and_edge_shapes():
    smooth, _ = _ref_smoothness()
    g = torch.Generator(device="cuda").manual_seed(1)
    for shape in ((1, 16, 150, 64), (1, 3, 3, 5), (2, 4, 7, 1), (1, 8, 64, 128)):
        t = torch.randn(shape, generator=g, device="cuda", requires_grad=True)
        r = smooth(t)
        r.backward()
        gr = t.grad.clone(); t.grad = None
        o = freg.compute_plane_smoothness(t)
        o.backward()
        assert abs(float(o) - float(r)) <= 2e-6 * abs(float(r)), shape
        assert util.rel_err(t.grad, gr) < 2e-6, shape


def _adam_pair(shapes, lrs, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    pa = [torch.nn.Parameter(torch.randn(s, generator=g, device="cuda")) for s in shapes]
    pb = [torch.nn.Parameter(p.detach().clone()) for p in pa]
    mk = lambda ps: [{"params": [p], "lr": lr, "name": f"g{i}"} for i, (p, lr) in enumerate(zip(ps, lrs))]
    return pa, pb, torch.optim.Adam(mk(pa), lr=0.0, eps=1e-15), fopt.FusedAdam(mk(pb), lr=0.0, eps=1e-15)


def test_fused_adam_bit_identical_to_torch_adam():
    shapes = [(30001, 3), (30001, 1, 3), (30001, 15, 3), (30001, 1), (30001, 4), (1, 16, 64, 64), (128, 32), (128,), (7,), (3, 128)]
    lrs = [1.6e-4, 2.5e-3, 1.25e-4, 5e-2, 1e-3, 1.6e-3, 1.6e-4, 1.6e-4, 1e-2, 0.0]
    pa, pb, ref, ours = _adam_pair(shapes, lrs)
    g = torch.Generator(device="cuda").manual_seed(9)
    for it in range(6):
        for i, (a, b) in enumerate(zip(pa, pb)):
            if i == 8 and it < 2:
                a.grad = None; b.grad = None          # a parameter that receives its first gradient late: own step count
                continue
            gr = torch.randn(a.shape, generator=g, device="cuda") * (10.0 ** (it - 3))
            if i == 0:
                gr[::3] = 0.0                         # invisible Gaussians: zero gradient, moments still decay
            a.grad = gr.clone(); b.grad = gr.clone()
        if it == 3:                                   # the reference's schedulers rewrite group lrs every iteration
            for opt in (ref, ours):
                opt.param_groups[0]["lr"] = 7.7e-5
        ref.step(); ours.step()
        for i, (a, b) in enumerate(zip(pa, pb)):
            assert torch.equal(a, b), (it, i, float((a - b).abs().max()))
            if a in ref.state:
                sa, sb = ref.state[a], ours.state[b]
                assert torch.equal(sa["exp_avg"], sb["exp_avg"]), (it, i)
                assert torch.equal(sa["exp_avg_sq"], sb["exp_avg_sq"]), (it, i)
                assert float(sa["step"]) == float(sb["step"])
    # state_dict layouts are interchangeable: load ours into a fresh torch Adam and keep stepping identically
    import copy
    sd = copy.deepcopy(ours.state_dict())        # load_state_dict keeps same-device tensors BY REFERENCE: without the copy
    assert set(sd["state"][0]) == {"step", "exp_avg", "exp_avg_sq"}      # both optimizers would update the same moments
    ref2 = torch.optim.Adam([{"params": [p], "lr": 0.0} for p in pa], lr=0.0, eps=1e-15)
    ref2.load_state_dict(sd)
    for a, b in zip(pa, pb):
        gr = torch.randn(a.shape, generator=g, device="cuda")
        a.grad = gr.clone(); b.grad = gr.clone()
    ref2.step(); ours.step(zero_grad=True)
    for a, b in zip(pa, pb):
        assert torch.equal(a, b)
        assert float(b.grad.abs().max()) == 0.0
    util.plog("train_glue", "FusedAdam: parameters, exp_avg, exp_avg_sq bit-identical to torch.optim.Adam over 7 steps, 10 tensors")


def test_fused_adam_errors():
    p = torch.nn.Parameter(torch.zeros(4))
    p.grad = torch.ones(4)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        fopt.FusedAdam([p], lr=1e-3).step()
    with pytest.raises(NotImplementedError):
        fopt.FusedAdam([p], lr=1e-3, weight_decay=0.1)


def test_densification_stats_vs_reference_expressions():
    P = 50_003
    g = torch.Generator(device="cuda").manual_seed(2)
    radii = torch.randint(0, 40, (P,), generator=g, device="cuda", dtype=torch.int32)
    radii[torch.rand(P, generator=g, device="cuda") < 0.3] = 0
    vs = torch.randn(P, 3, generator=g, device="cuda")
    max_r = torch.rand(P, generator=g, device="cuda") * 30
    accum = torch.rand(P, 1, generator=g, device="cuda")
    denom = torch.randint(0, 5, (P, 1), generator=g, device="cuda").float()
    # reference: train.py:346-349, gaussian_model.py:521-523
    vis = radii > 0
    r_max, r_acc, r_den = max_r.clone(), accum.clone(), denom.clone()
    r_max[vis] = torch.max(r_max[vis], radii[vis])
    r_acc[vis] += torch.norm(vs[vis, :2], dim=-1, keepdim=True)
    r_den[vis] += 1
    fopt.add_densification_stats(radii, vs, max_r, accum, denom)
    assert torch.equal(max_r, r_max) and torch.equal(denom, r_den)
    assert util.rel_err(accum, r_acc) < 2e-7
    # explicit visibility mask (a subset of radii > 0)
    vis2 = vis & (torch.rand(P, generator=g, device="cuda") < 0.5)
    m2, a2, d2 = r_max.clone(), r_acc.clone(), r_den.clone()
    m2[vis2] = torch.max(m2[vis2], radii[vis2]); a2[vis2] += torch.norm(vs[vis2, :2], dim=-1, keepdim=True); d2[vis2] += 1
    fopt.add_densification_stats(radii, vs, max_r, accum, denom, visibility=vis2)
    assert torch.equal(max_r, m2) and torch.equal(denom, d2) and util.rel_err(accum, a2) < 2e-7
