"""GPU parity of the sm_100a rasterizer (through the C ABI and the drop-in GaussianRasterizer) against
  (a) golden fixtures produced by the UNMODIFIED reference CUDA build (tests/golden/raster_*.npz),
  (b) the CPU restatement (oracle/raster_oracle.cpp), FP32-with-FMA-emulation and FP64 arbiter,
  (c) the reference build itself when oracle/_ref travelled to this box,
and through size-independent properties at the full BASELINE sizes.

Difference categories (BASELINE.json north_star):
  sort keys, point list, tile ranges, radii, tiles_touched, n_contrib : BIT-EXACT
  depths / means2D / conic / cov3D / final_T / depth image             : bit-exact vs the reference build
  colour image                                                         : <= 1e-5 relative
  gradients: the reference's float atomics are nondeterministic; tolerance = max(20 x its recorded
  run-to-run spread, 2e-5) relative to the tensor's max magnitude.
"""
import os

import numpy as np
import pytest
import torch

import fdgs  # noqa: F401
import util

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
IMG_TOL = 1e-5
CASES = ["tiny", "ragged", "dense", "deg0", "precomp"]


def _golden_scene(name):
    z = np.load(os.path.join(GOLD, f"raster_{name}.npz"))
    sc = dict(W=int(z["meta/W"]), H=int(z["meta/H"]), tanfovx=float(z["meta/tanfovx"]), tanfovy=float(z["meta/tanfovy"]),
              sh_degree=int(z["meta/sh_degree"]))
    for k in z.files:
        if k.startswith("in/") and k != "in/dL_dpix":
            sc[k[3:]] = torch.tensor(z[k]).cuda()
    return z, sc, float(z["meta/scale_modifier"])


def _eq(a, b, what):
    a = a.detach().cpu().numpy() if isinstance(a, torch.Tensor) else np.asarray(a)
    assert a.shape == tuple(b.shape), what
    assert np.array_equal(a.view(np.uint8) if a.dtype.kind == "f" else a, b.view(np.uint8) if b.dtype.kind == "f" else b), what


def _have_golden():
    return os.path.exists(os.path.join(GOLD, "raster_tiny.npz"))


@pytest.mark.skipif(not _have_golden(), reason="raster golden fixtures not generated yet")
@pytest.mark.parametrize("name", CASES)
def test_forward_vs_reference_golden(name):
    z, sc, smod = _golden_scene(name)
    o = util.ours_forward(sc, scale_modifier=smod)
    assert o["num_rendered"] == int(z["meta/num_rendered"])
    _eq(o["radii"], z["out/radii"], "radii")
    vis = torch.tensor(z["out/radii"] > 0).cuda()
    _eq(o["tiles_touched"] * vis, z["out/tiles_touched"], "tiles_touched")
    _eq(o["keys"], z["out/keys"], "keys")
    _eq(o["point_list"], z["out/point_list"], "point_list")
    _eq(o["ranges"], z["out/ranges"], "ranges")
    _eq(o["n_contrib"], z["out/n_contrib"], "n_contrib")
    for k in ("depths", "means2D", "conic_opacity", "cov3D"):
        if k == "cov3D" and sc.get("cov3D_precomp") is not None:
            continue     # the reference does not write cov3D when it is supplied
        m = vis if o[k].dim() == 1 else vis[:, None].expand_as(o[k])
        _eq(torch.where(m, o[k], torch.zeros_like(o[k])), z["out/" + k], k)
    _eq(o["final_T"], z["out/final_T"], "final_T")
    _eq(o["depth"], z["out/depth"], "depth image")
    assert util.rel_err(o["color"], z["out/color"]) < IMG_TOL


@pytest.mark.skipif(not _have_golden(), reason="raster golden fixtures not generated yet")
@pytest.mark.parametrize("name", CASES)
def test_backward_vs_reference_golden(name):
    z, sc, smod = _golden_scene(name)
    gc = torch.tensor(z["in/dL_dpix"]).cuda()
    color, radii, depth, g = util.ours_autograd(sc, gc, scale_modifier=smod)
    pairs = [("means3D", "means3D"), ("means2D", "means2D"), ("opacities", "opacities")]
    if sc.get("shs") is not None:
        pairs += [("shs", "sh")]
    else:
        pairs += [("colors_precomp", "colors_precomp")]
    if sc.get("scales") is not None:
        pairs += [("scales", "scales"), ("rotations", "rotations")]
    else:
        pairs += [("cov3D_precomp", "cov3D")]
    for ours_k, ref_k in pairs:
        ref = z["g/" + ref_k]
        tol = max(20.0 * float(z["gspread/" + ref_k]), 2e-5)
        a = g[ours_k]
        assert a is not None, ours_k
        assert a.numel() == ref.size, ours_k
        assert util.rel_err(a.reshape(-1), ref.reshape(-1)) < tol, (ours_k, tol)
    assert float(g["means2D"][:, 2].abs().max()) == 0.0          # .z is never written (backward.cu:545-546)


@pytest.mark.parametrize("kw", [dict(P=2000, W=200, H=120, logscale=-3.0, seed=3),
                                dict(P=4096, W=333, H=217, logscale=-2.5, seed=9, bg=0.0),
                                dict(P=500, W=256, H=256, logscale=-0.5, seed=4),
                                dict(P=257, W=31, H=17, logscale=-2.0, seed=1, sh_degree=2)])
def test_forward_vs_cpu_oracle(kw):
    sc = util.scene(**kw)
    scd = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in sc.items()}
    o = util.ours_forward(scd)
    c = util.oracle_forward(sc)
    assert o["num_rendered"] == c["num_rendered"]
    _eq(o["radii"], c["radii"], "radii")
    _eq(o["keys"].cpu().numpy().astype(np.uint64), c["keys"], "keys")
    _eq(o["point_list"].cpu().numpy().astype(np.uint32), c["point_list"], "point_list")
    _eq(o["ranges"].cpu().numpy().astype(np.uint32), c["ranges"], "ranges")
    # the oracle's libm expf differs from CUDA's by an ulp, which can move a pixel across the 1/255 or 1e-4
    # thresholds: allow a handful of pixels to stop one contributor earlier/later
    nc = o["n_contrib"].cpu().numpy().astype(np.int64)
    bad = nc != c["n_contrib"].astype(np.int64)
    assert bad.sum() <= max(2, nc.size // 20000)
    ok = torch.tensor(~bad)
    # the same ulp can move one contributor across the alpha < 1/255 skip (forward.cu:336) without changing
    # n_contrib: such a pixel differs by <= alpha*|colour| ~ 4e-3. Allow the same handful of pixels, bound their
    # error by that skip size, and hold every other pixel to IMG_TOL.
    for name in ("color", "depth"):
        a = o[name].cpu().double()[:, ok]
        b = torch.tensor(c[name]).double()[:, ok]
        err = ((a - b).abs() / (b.abs().max() + 1e-30)).max(0).values
        flips = err >= IMG_TOL
        assert int(flips.sum()) <= max(2, nc.size // 20000), (name, int(flips.sum()), float(err.max()))
        assert float(err.max()) < 5e-3, (name, float(err.max()))


@pytest.mark.parametrize("kw", [dict(P=1200, W=160, H=96, logscale=-2.6, seed=3), dict(P=300, W=64, H=64, logscale=-1.0, seed=8, bg=0.0)])
def test_backward_vs_fp64_oracle(kw):
    sc = util.scene(**kw)
    scd = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in sc.items()}
    gen = torch.Generator().manual_seed(5)
    gc = torch.randn((3, sc["H"], sc["W"]), generator=gen)
    _, _, _, g = util.ours_autograd(scd, gc.cuda())
    c = util.oracle_forward(sc, dL_dpix=gc, f64=True)
    for k, ck in (("means3D", "dmeans3D"), ("means2D", "dmean2D"), ("opacities", "dopacity"), ("scales", "dscale"),
                  ("rotations", "drot"), ("shs", "dsh")):
        # FP32 kernel vs FP64 arbiter: 1e-3 of the tensor's max (threshold pixels may flip in FP64)
        assert util.rel_err(g[k].reshape(-1), c[ck].reshape(-1)) < 1e-3, k


def test_determinism_forward_and_backward():
    """Two runs are bit-identical in forward; backward is reduced per (tile, Gaussian) in a fixed order
    inside a tile but tiles still meet in global atomics, so only forward is asserted bit-exact and the
    backward must agree to 1e-6."""
    sc = util.scene(P=3000, W=256, H=192, logscale=-2.8, seed=12)
    scd = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in sc.items()}
    a, b = util.ours_forward(scd), util.ours_forward(scd)
    for k in ("keys", "point_list", "ranges", "n_contrib", "final_T", "color", "depth", "radii"):
        _eq(a[k], b[k].cpu().numpy(), k)
    gc = torch.randn((3, 192, 256), generator=torch.Generator().manual_seed(1)).cuda()
    g1, g2 = util.ours_autograd(scd, gc)[3], util.ours_autograd(scd, gc)[3]
    for k in g1:
        if g1[k] is not None:
            assert util.rel_err(g1[k], g2[k]) < 1e-6, k


def _settings(sc):
    return util.settings(sc)


def test_edge_empty_and_all_culled():
    from fdgs.diff_gaussian_rasterization import GaussianRasterizer
    sc = util.scene(P=64, W=50, H=34, logscale=-2.0, seed=2, device="cuda")
    rast = GaussianRasterizer(_settings(sc))
    z = lambda *s: torch.zeros(s, device="cuda")
    # P == 0: zeros, like rasterize_points.cu:82
    color, radii, depth = rast(means3D=z(0, 3), means2D=z(0, 3), opacities=z(0, 1), shs=z(0, 16, 3), scales=z(0, 3), rotations=z(0, 4))
    assert color.shape == (3, 34, 50) and float(color.abs().max()) == 0 and radii.numel() == 0 and depth.shape == (1, 34, 50)
    # everything behind the camera: R == 0 -> background everywhere, final_T == 1
    behind = sc["campos"][None, :] * 3.0 + torch.randn(64, 3, device="cuda") * 0.01   # camera looks at the origin
    o = util.ours_forward(dict(sc, means3D=behind))
    assert o["num_rendered"] == 0 and int(o["radii"].abs().max()) == 0
    assert float((o["color"] - 1.0).abs().max()) == 0 and float(o["depth"].abs().max()) == 0
    assert float((o["final_T"] - 1).abs().max()) == 0 and int(o["n_contrib"].max()) == 0
    # and its backward is all zeros
    _, _, _, g = util.ours_autograd(dict(sc, means3D=behind), torch.ones(3, 34, 50, device="cuda"))
    for k, v in g.items():
        if v is not None:
            assert float(v.abs().max()) == 0, k


def test_edge_screen_filling_gaussian_and_equal_depths():
    """One Gaussian covering every tile (warp-cooperative key emission) + exact depth ties (stable order = id)."""
    sc = util.scene(P=40, W=130, H=70, logscale=-2.0, seed=7)
    sc["scales"][0] = torch.tensor([3.0, 3.0, 3.0]); sc["means3D"][0] = torch.tensor([0.0, 0.0, 0.0])
    sc["means3D"][10:20] = sc["means3D"][10].clone()          # identical positions -> identical depth bits
    scd = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in sc.items()}
    o = util.ours_forward(scd)
    c = util.oracle_forward(sc)
    T = ((130 + 15) // 16) * ((70 + 15) // 16)
    assert int(o["tiles_touched"][0]) == T
    _eq(o["keys"].cpu().numpy().astype(np.uint64), c["keys"], "keys")
    _eq(o["point_list"].cpu().numpy().astype(np.uint32), c["point_list"], "point_list")
    keys = o["keys"].cpu().numpy().astype(np.uint64)
    assert np.all(keys[1:] >= keys[:-1])
    pl = o["point_list"].cpu().numpy()
    same = keys[1:] == keys[:-1]
    assert same.any() and np.all(pl[1:][same] > pl[:-1][same])


def test_sort_and_range_properties_full_size():
    """BASELINE config 4 size (300k Gaussians, 1352x1014): sortedness, stable tie order, ranges == searchsorted
    on the tile ids, sum(tiles_touched) == R, and agreement with numpy's stable argsort of the emitted keys."""
    sc = util.scene(P=300_000, W=1352, H=1014, fovx=1.2, extent=2.0, logscale=-4.828, seed=0, device="cuda")
    o = util.ours_forward(sc)
    R = o["num_rendered"]
    keys = o["keys"].cpu().numpy().astype(np.uint64)
    pl = o["point_list"].cpu().numpy().astype(np.int64)
    assert R == keys.size == int(o["tiles_touched"].sum())
    assert np.all(keys[1:] >= keys[:-1])
    same = keys[1:] == keys[:-1]
    assert np.all(pl[1:][same] > pl[:-1][same])
    tiles = (keys >> np.uint64(32)).astype(np.int64)
    T = ((1352 + 15) // 16) * ((1014 + 15) // 16)
    rg = o["ranges"].cpu().numpy().astype(np.int64)
    lo = np.searchsorted(tiles, np.arange(T), side="left"); hi = np.searchsorted(tiles, np.arange(T), side="right")
    nonempty = hi > lo
    assert np.array_equal(rg[nonempty, 0], lo[nonempty]) and np.array_equal(rg[nonempty, 1], hi[nonempty])
    assert not rg[~nonempty].any()
    # depth bits of the key are the Gaussian's view depth
    d = o["depths"].cpu().numpy().view(np.uint32)
    assert np.array_equal((keys & np.uint64(0xFFFFFFFF)).astype(np.uint32), d[pl])
    # image sanity: T in [0,1], contributors within the tile's list
    ft = o["final_T"]
    assert float(ft.min()) >= 0 and float(ft.max()) <= 1
    assert torch.isfinite(o["color"]).all() and torch.isfinite(o["depth"]).all()


def test_mark_visible():
    from fdgs.diff_gaussian_rasterization import GaussianRasterizer
    sc = util.scene(P=5000, W=64, H=64, seed=3, device="cuda", dist=1.0)
    vis = GaussianRasterizer(_settings(sc)).markVisible(sc["means3D"])
    pv = torch.cat([sc["means3D"], torch.ones(5000, 1, device="cuda")], 1) @ sc["viewmatrix"]
    assert vis.dtype == torch.bool and torch.equal(vis, pv[:, 2] > 0.2)


def test_api_validation_and_errors():
    from fdgs.diff_gaussian_rasterization import GaussianRasterizer
    sc = util.scene(P=16, W=32, H=32, seed=3, device="cuda")
    rast = GaussianRasterizer(_settings(sc))
    m2 = torch.zeros_like(sc["means3D"])
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        rast(means3D=sc["means3D"], means2D=m2, opacities=sc["opacities"], scales=sc["scales"], rotations=sc["rotations"])
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        rast(means3D=sc["means3D"], means2D=m2, opacities=sc["opacities"], shs=sc["shs"])
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        rast(means3D=sc["means3D"].cpu(), means2D=m2.cpu(), opacities=sc["opacities"].cpu(), shs=sc["shs"].cpu(),
             scales=sc["scales"].cpu(), rotations=sc["rotations"].cpu())
    with pytest.raises(RuntimeError, match="num_points, 3"):
        rast(means3D=sc["means3D"][:, :2], means2D=m2, opacities=sc["opacities"], shs=sc["shs"], scales=sc["scales"],
             rotations=sc["rotations"])
    # [1,4,4]-shaped camera matrices (scene/dataset_readers.py:522-546) are accepted
    s2 = _settings(sc)._replace(viewmatrix=sc["viewmatrix"][None], projmatrix=sc["projmatrix"][None], debug=True)
    c2, _, _ = GaussianRasterizer(s2)(means3D=sc["means3D"], means2D=m2, opacities=sc["opacities"], shs=sc["shs"],
                                      scales=sc["scales"], rotations=sc["rotations"])
    c1, _, _ = rast(means3D=sc["means3D"], means2D=m2, opacities=sc["opacities"], shs=sc["shs"],
                    scales=sc["scales"], rotations=sc["rotations"])
    assert torch.equal(c1, c2)


def _ref_or_skip():
    import build_ref
    if not build_ref.available():
        pytest.skip("oracle/_ref not built")
    return build_ref.load()


BIT_EXACT = ("radii", "keys", "point_list", "ranges", "n_contrib", "final_T", "depth")


def _compare_with_reference_build(refC, sc, tag, backward=True):
    o, r = util.ours_forward(sc), util.ref_forward(refC, sc)
    assert o["num_rendered"] == r["num_rendered"]
    for k in BIT_EXACT:
        assert torch.equal(o[k].reshape(-1), r[k].reshape(-1)), k
    vis = r["radii"] > 0
    for k in ("depths", "means2D", "conic_opacity", "cov3D"):         # per-Gaussian geometry of the visible ones: bit-exact
        assert torch.equal(o[k][vis], r[k][vis]), k
    e_img = util.rel_err(o["color"], r["color"])
    assert e_img < IMG_TOL
    maxb = int((o["ranges"][:, 1] - o["ranges"][:, 0]).max())
    line = f"{tag}: R={o['num_rendered']} max_bucket={maxb} bit-exact {','.join(BIT_EXACT)} + geometry; colour rel err {e_img:.2e}"
    if backward:
        gc = torch.randn(o["color"].shape, generator=torch.Generator().manual_seed(5)).cuda()
        g = util.ours_autograd(sc, gc)[3]
        g1, g2 = util.ref_backward(refC, sc, r, gc), util.ref_backward(refC, sc, r, gc)
        for k, rk in (("means3D", "means3D"), ("means2D", "means2D"), ("opacities", "opacities"), ("scales", "scales"),
                      ("rotations", "rotations"), ("shs", "sh")):
            spread = util.rel_err(g1[rk], g2[rk])
            tol = max(20 * spread, 2e-5)
            e = util.rel_err(g[k].reshape(-1), g1[rk].reshape(-1))
            line += f"; d{k} {e:.1e} (ref spread {spread:.1e})"
            assert e < tol, k
    util.plog("raster", line)
    return o


@pytest.mark.parametrize("kw", [dict(P=50_000, W=800, H=800, logscale=-4.605, seed=0),
                                dict(P=4096, W=333, H=217, logscale=-2.5, seed=9, bg=0.0)])
def test_vs_reference_build_when_present(kw):
    _compare_with_reference_build(_ref_or_skip(), util.scene(device="cuda", **kw), f"ref_build[P={kw['P']}]")


# BASELINE.json configs 2-5 (SURVEY 8d shapes), one view each: "sort keys, tile ranges and radii are bit-exact"
BASELINE_RASTER = {
    "cfg2_dnerf_200k": dict(P=200_000, W=800, H=800, fovx=0.6911, extent=1.3, logscale=-4.828, seed=0),
    "cfg3_hypernerf_300k": dict(P=300_000, W=536, H=960, fovx=0.9, extent=2.0, logscale=-4.828, seed=0),
    "cfg4_dynerf_300k": dict(P=300_000, W=1352, H=1014, fovx=1.2, extent=2.0, logscale=-4.828, seed=0),
    "cfg5_infer_1M": dict(P=1_000_000, W=1920, H=1080, fovx=1.2, extent=2.0, logscale=-5.298, seed=0),
}


@pytest.mark.parametrize("name", list(BASELINE_RASTER))
def test_bit_exact_vs_reference_build_at_baseline_configs(name):
    refC = _ref_or_skip()
    sc = util.scene(device="cuda", **BASELINE_RASTER[name])
    _compare_with_reference_build(refC, sc, name, backward=(name != "cfg5_infer_1M"))     # config 5 is render-only


# Clustered scenes: every Gaussian lands in a handful of tiles, so ONE tile bucket holds thousands of instances — the
# load-imbalance case of real scenes that the uniform clouds above (mean ~350 / tile) never produce.  The
# sizes straddle every bucket-size class of the segmented sort (<= 2048, <= 8192, <= 28672, larger).
@pytest.mark.parametrize("P,extent", [(1500, 0.03), (6000, 0.03), (20_000, 0.03), (45_000, 0.03), (30_000, 0.3)])
def test_clustered_scene_large_tile_buckets_vs_reference_build(P, extent):
    refC = _ref_or_skip()
    sc = util.scene(P=P, W=256, H=160, extent=extent, logscale=-5.5, seed=21, device="cuda")
    o = _compare_with_reference_build(refC, sc, f"clustered[P={P},extent={extent}]")
    maxb = int((o["ranges"][:, 1] - o["ranges"][:, 0]).max())
    if extent < 0.1:
        assert maxb > P // 8, maxb          # the scene really is clustered: one tile holds a large share of all instances


def test_clustered_scene_vs_cpu_oracle():
    """Same large-bucket path without the reference build on the box: keys / point list / ranges vs the CPU restatement."""
    sc = util.scene(P=9000, W=128, H=96, extent=0.03, logscale=-5.5, seed=22)
    scd = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in sc.items()}
    o = util.ours_forward(scd)
    c = util.oracle_forward(sc)
    assert o["num_rendered"] == c["num_rendered"]
    assert int((o["ranges"][:, 1] - o["ranges"][:, 0]).max()) > 2048
    _eq(o["keys"].cpu().numpy().astype(np.uint64), c["keys"], "keys")
    _eq(o["point_list"].cpu().numpy().astype(np.uint32), c["point_list"], "point_list")
    _eq(o["ranges"].cpu().numpy().astype(np.uint32), c["ranges"], "ranges")


def test_one_shot_forward_matches_exact_and_reports_overflow():
    """fdgs_raster_forward (no host sync; workspace sized from an estimate) vs the plan / render pair: identical images and
    gradients when the estimate holds; when it does not, nothing is rendered, check_pending() says so, and the exact path
    afterwards is unaffected."""
    from fdgs import diff_gaussian_rasterization as R
    sc = util.scene(P=30_000, W=400, H=300, logscale=-4.0, seed=5, device="cuda")
    gc = torch.randn((3, 300, 400), generator=torch.Generator().manual_seed(1)).cuda()
    assert R.check_pending()
    c0, r0, d0, g0 = util.ours_autograd(sc, gc)                      # exact path; also records the count for this shape
    key = (torch.device("cuda", torch.cuda.current_device()).index, 30_000, 400, 300)
    assert R._caps[key] == R.last_num_rendered() > 0
    with R.speculative_mode(True):
        c1, r1, d1, g1 = util.ours_autograd(sc, gc)
        assert len(R._pending) == 1
    assert R.check_pending(block=True) and not R._pending
    assert torch.equal(c0, c1) and torch.equal(r0, r1) and torch.equal(d0, d1)
    for k in g0:
        if g0[k] is not None:
            assert util.rel_err(g1[k], g0[k]) < 5e-6, k             # same kernels; float atomics in the backward (run-to-run noise)
    # overflow: pretend the last render of this shape was 20x smaller
    true_R = R._caps[key]
    R._caps[key] = true_R // 20
    R.SPEC_SLACK, slack = 0, R.SPEC_SLACK
    try:
        with R.speculative_mode(True):
            c2, r2, d2, g2 = util.ours_autograd(sc, gc)
        assert not R.check_pending(block=True)                        # reported ...
        assert float((c2 - 1.0).abs().max()) == 0.0                   # ... nothing was blended (background = 1)
        assert torch.equal(r2, r0)                                    # radii come from the preprocess stage
        assert all(float(v.abs().max()) == 0.0 for k, v in g2.items() if v is not None and k != "means2D")
        assert R._caps[key] >= true_R                                 # the real count was learnt from the failed call
        with R.speculative_mode(True):
            c3, _, _, _ = util.ours_autograd(sc, gc)
        assert R.check_pending(block=True) and torch.equal(c3, c0)
    finally:
        R.SPEC_SLACK = slack
