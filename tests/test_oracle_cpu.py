"""CPU (`-m "not gpu"`) pinning of the two oracles against fixtures produced by the REAL reference:

* tests/golden/deform_*.npz  — outputs and gradients of the reference `deform_network` module itself
  (oracle/make_golden.py imports scene/deformation.py + scene/hexplane.py from /root/reference);
* tests/golden/raster_*.npz  — outputs, decoded workspaces and gradients of the UNMODIFIED reference CUDA
  rasterizer built for sm_100 and run on a B200 (tests/golden/make_raster_golden.py).

The restatements (oracle/deform_oracle.py, oracle/raster_oracle.cpp) must reproduce them: integers bit-exact,
floats within the tolerances written below.  Nothing here touches the CUDA product path."""
import os

import numpy as np
import pytest
import torch

import fdgs  # noqa: F401
from fdgs.configs import hidden_params, _kp
import deform_oracle as DO
import raster_oracle as RO

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

DEFORM_CASES = {
    "dynerf128": dict(kplanes_config=_kp(16, [16, 16, 16, 20]), multires=[1, 2], defor_depth=0, net_width=128, no_do=False,
                      no_dshs=False),
    "tiny": dict(kplanes_config=_kp(8, [16, 16, 16, 20]), multires=[1], defor_depth=0, net_width=32, no_do=False, no_dshs=False),
    "small3": dict(kplanes_config=_kp(16, [12, 10, 8, 9]), multires=[1, 2], defor_depth=1, net_width=64),
    "bcast": dict(kplanes_config=_kp(8, [8, 8, 8, 6]), multires=[1], defor_depth=0, net_width=32, no_do=False, no_dshs=False),
}


def _rel(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / (np.abs(b).max() + 1e-30))


@pytest.mark.parametrize("name", list(DEFORM_CASES))
@pytest.mark.parametrize("dtype,out_tol,grad_tol", [(torch.float32, 2e-6, 2e-5), (torch.float64, 5e-7, 5e-6)])
def test_deform_oracle_matches_reference_module(name, dtype, out_tol, grad_tol):
    """FP32 restatement vs the reference's FP32 results: summation-order noise only.  The FP64 copy (the
    arbiter used by the GPU tests) must sit within FP32 rounding of the same fixtures."""
    z = np.load(os.path.join(GOLD, f"deform_{name}.npz"))
    if name == "dynerf128" and dtype == torch.float64:
        # 2001 points x five 128-wide heads: the FP32 fixture's own summation error on a plane gradient is 5.5e-6 of
        # the largest entry (measured against this FP64 copy), above the 5e-6 the smaller cases need
        grad_tol = 1e-5
    cfg = DO.cfg_from_args(hidden_params(**DEFORM_CASES[name]))
    sd = {k[3:]: torch.tensor(z[k]).to(dtype).requires_grad_(k[3:].startswith("deformation_net.") and "aabb" not in k)
          for k in z.files if k.startswith("sd/")}
    inp = {k[3:]: torch.tensor(z[k]).to(dtype).requires_grad_(True) for k in z.files if k.startswith("in/")}
    P = inp["xyz"].shape[0]
    t = torch.full((P, 1), float(z["meta/t"]), dtype=dtype)
    outs = DO.deform_forward(sd, cfg, inp["xyz"], inp["scales"], inp["rot"], inp["opacity"], inp["shs"], t)
    loss = sum((o * torch.tensor(z[f"w/{i}"]).to(dtype)).sum() for i, o in enumerate(outs))
    loss.backward()
    for i, o in enumerate(outs):
        assert tuple(o.shape) == z[f"out/{i}"].shape
        assert _rel(o.detach().numpy(), z[f"out/{i}"]) < out_tol, f"out[{i}]"
    for k, v in inp.items():
        ref = z["gin/" + k]
        g = v.grad
        if g is None:
            assert not np.any(ref), k
        else:
            assert _rel(g.numpy(), ref) < grad_tol, k
    n = 0
    for k in z.files:
        if k.startswith("gp/"):
            g = sd[k[3:]].grad
            assert g is not None, k
            assert _rel(g.numpy(), z[k]) < grad_tol, k
            n += 1
    assert n >= 14


@pytest.mark.parametrize("name", list(DEFORM_CASES))
def test_ref_port_is_bit_identical_to_the_reference_module(name):
    """oracle/ref_port.py (fallback reference arm / GPU cross-check) issues the SAME ATen ops as the reference module:
    on CPU its outputs and gradients equal the fixtures the real module produced BIT FOR BIT; and when the real
    module is loadable (authoring container, or baseline/_ref on the box) it is compared directly as well."""
    import ref_port
    z = np.load(os.path.join(GOLD, f"deform_{name}.npz"))
    args = hidden_params(**DEFORM_CASES[name])
    sd = {k[3:]: torch.tensor(z[k]) for k in z.files if k.startswith("sd/")}

    def run(make):
        net = make(args)
        net.deformation_net.set_aabb(sd["deformation_net.grid.aabb"][0].tolist(), sd["deformation_net.grid.aabb"][1].tolist())
        net.load_state_dict(sd)
        inp = {k[3:]: torch.tensor(z[k]).requires_grad_(True) for k in z.files if k.startswith("in/")}
        P = inp["xyz"].shape[0]
        outs = net(inp["xyz"], inp["scales"], inp["rot"], inp["opacity"], inp["shs"], torch.full((P, 1), float(z["meta/t"])))
        sum((o * torch.tensor(z[f"w/{i}"])).sum() for i, o in enumerate(outs)).backward()
        return outs, inp, dict(net.named_parameters())

    outs, inp, params = run(ref_port.ref_deform_network)
    for i, o in enumerate(outs):
        assert np.array_equal(o.detach().numpy(), z[f"out/{i}"]), f"out[{i}]"
    for k, v in inp.items():
        assert np.array_equal(v.grad.numpy() if v.grad is not None else np.zeros_like(z["gin/" + k]), z["gin/" + k]), k
    for k in z.files:
        if k.startswith("gp/"):
            assert np.array_equal(params[k[3:]].grad.numpy(), z[k]), k
    import ref_deform
    if ref_deform.available():
        outs_r, inp_r, params_r = run(ref_deform.make_network)
        for a, b in zip(outs, outs_r):
            assert torch.equal(a, b)
        for k in params_r:
            ga, gb = params[k].grad, params_r[k].grad
            assert (ga is None and gb is None) or torch.equal(ga, gb), k


RASTER_CASES = ["tiny", "deg0", "dense", "precomp", "ragged"]


def _raster_inputs(z):
    g = lambda k: z[k] if k in z.files else None
    return dict(W=int(z["meta/W"]), H=int(z["meta/H"]), tanfovx=float(z["meta/tanfovx"]), tanfovy=float(z["meta/tanfovy"]),
                bg=z["in/bg"], means3D=z["in/means3D"], opacities=z["in/opacities"], viewmatrix=z["in/viewmatrix"],
                projmatrix=z["in/projmatrix"], campos=z["in/campos"], shs=g("in/shs"), colors_precomp=g("in/colors_precomp"),
                scales=g("in/scales"), rotations=g("in/rotations"), cov3D_precomp=g("in/cov3D_precomp"),
                sh_degree=int(z["meta/sh_degree"]), scale_modifier=float(z["meta/scale_modifier"]))


@pytest.mark.parametrize("name", RASTER_CASES)
def test_raster_oracle_forward_matches_reference_build(name):
    """Keys, sorted Gaussian ids, tile ranges, radii, tiles_touched: bit-exact.  Per-Gaussian floats: the depth
    bits are exact (they are the low half of the keys); images / depth within 1e-5 of the image maximum."""
    z = np.load(os.path.join(GOLD, f"raster_{name}.npz"))
    o = RO.rasterize(**_raster_inputs(z))
    assert o["num_rendered"] == int(z["meta/num_rendered"])
    assert np.array_equal(o["radii"], z["out/radii"])
    assert np.array_equal(o["tiles_touched"].astype(np.int64), z["out/tiles_touched"].astype(np.int64))
    assert np.array_equal(o["keys"], z["out/keys"].astype(np.uint64))
    assert np.array_equal(o["point_list"], z["out/point_list"].astype(np.uint32))
    assert np.array_equal(o["ranges"], z["out/ranges"].astype(np.uint32))
    vis = z["out/radii"] > 0
    assert np.array_equal(o["depths"][vis].view(np.uint32), z["out/depths"][vis].view(np.uint32))
    assert _rel(o["means2D"][vis], z["out/means2D"][vis]) < 1e-6
    assert _rel(o["conic_opacity"][vis], z["out/conic_opacity"][vis]) < 1e-5
    if "in/colors_precomp" not in z.files:     # with precomputed colours the reference never writes its rgb scratch
        assert _rel(o["rgb"][vis], z["out/rgb"][vis]) < 1e-6
    nc_bad = o["n_contrib"].astype(np.int64) != z["out/n_contrib"].astype(np.int64)
    assert nc_bad.sum() <= 2          # libm expf vs CUDA expf: a threshold pixel may stop one contributor apart
    ok = ~nc_bad
    assert _rel(o["color"][:, ok], z["out/color"][:, ok]) < 1e-5
    assert _rel(o["depth"][:, ok], z["out/depth"][:, ok]) < 1e-5
    assert _rel(o["final_T"][ok], z["out/final_T"][ok]) < 1e-5


@pytest.mark.parametrize("name", RASTER_CASES)
def test_raster_oracle_backward_matches_reference_build(name):
    """FP64 arbiter vs the reference's FP32 atomics: within 1e-4 of each tensor's maximum, and never worse than
    20x the reference's own run-to-run spread where that is larger (gspread/* in the fixture)."""
    z = np.load(os.path.join(GOLD, f"raster_{name}.npz"))
    o = RO.rasterize(**_raster_inputs(z), dL_dpix=z["in/dL_dpix"], f64=True)
    pairs = [("means2D", "dmean2D"), ("opacities", "dopacity"), ("means3D", "dmeans3D")]
    if "in/colors_precomp" in z.files:
        pairs.append(("colors_precomp", "dcolor"))
    else:
        pairs.append(("sh", "dsh"))
    if "in/scales" in z.files:
        pairs += [("scales", "dscale"), ("rotations", "drot")]
    else:
        pairs.append(("cov3D", "dcov3D"))
    for gk, ok in pairs:
        ref = z["g/" + gk]
        tol = max(1e-4, 20.0 * float(z["gspread/" + gk]))
        got = o[ok].reshape(ref.shape) if o[ok].size == ref.size else o[ok][..., :ref.shape[-1]]
        assert _rel(got, ref) < tol, (gk, _rel(got, ref), tol)


def test_raster_oracle_empty_and_culled_inputs():
    """P == 0 and an all-culled scene (everything behind the camera) render the background and emit no keys
    (rasterize_points.cu:82, rasterizer_impl.cu:314)."""
    z = np.load(os.path.join(GOLD, "raster_tiny.npz"))
    a = _raster_inputs(z)
    b = dict(a)
    b["means3D"] = a["means3D"] * 0 + np.array([0, 0, 50.0], np.float32)      # behind the +z camera looking at origin
    o = RO.rasterize(**b)
    assert o["num_rendered"] == 0 and not o["radii"].any()
    assert np.allclose(o["color"], a["bg"][:, None, None])
    e = dict(a)
    for k in ("means3D", "opacities", "shs", "scales", "rotations"):
        e[k] = a[k][:0]
    o = RO.rasterize(**e)
    assert o["num_rendered"] == 0
    assert np.allclose(o["color"], a["bg"][:, None, None])


def test_init_oracle_knn_against_kdtree_and_voxel_counts():
    """oracle/init_oracle.py: the brute-force 3-NN mean squared distance equals scipy's exact k-d tree query (float64, so
    equal up to FP32 rounding of the restated expression), and the voxel down-sampling conserves mass."""
    import init_oracle
    from scipy.spatial import cKDTree
    rng = np.random.default_rng(3)
    p = rng.uniform(-1, 1, size=(3000, 3)).astype(np.float32)
    got = init_oracle.knn_mean_dist2(p)
    d, _ = cKDTree(p.astype(np.float64)).query(p.astype(np.float64), k=4)
    want = (d[:, 1:] ** 2).mean(axis=1)
    assert np.allclose(got, want, rtol=2e-6, atol=0)
    c = rng.uniform(0, 1, size=(3000, 3)).astype(np.float32)
    vp, vc = init_oracle.voxel_downsample(p, c, 0.25)
    assert 1 < vp.shape[0] < 3000 and vc.shape == vp.shape
    one, _ = init_oracle.voxel_downsample(p, c, 100.0)          # one voxel: the centroid
    assert one.shape == (1, 3) and np.allclose(one[0], p.astype(np.float64).mean(0), atol=1e-6)
