"""TEST INFRASTRUCTURE — generates tests/golden/raster_*.npz by running the UNMODIFIED reference CUDA
rasterizer (oracle/_ref, built from /root/reference's own sources by oracle/build_ref.py) on seeded
inputs ON THE GPU BOX.  The fixtures hold the inputs and everything the parity tests compare: radii,
depths, sorted keys / point_list, tile ranges, n_contrib, final_T, colour, depth and the gradients of a
seeded dL/dcolour.  Backward is run 3 times; `gspread/*` records the reference's own run-to-run
spread (its atomics are nondeterministic), from which the gradient tolerance is derived.

Usage (GPU box):  python tests/golden/make_raster_golden.py     -> writes into gpurun_out/golden/
then copy the files into tests/golden/ and commit them.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import util  # noqa: E402
import build_ref  # noqa: E402

CASES = {
    # name: scene kwargs (+ variants)
    "tiny": dict(P=400, W=96, H=64, logscale=-2.5, seed=3, sh_degree=3, bg=1.0),
    "ragged": dict(P=600, W=83, H=61, logscale=-2.2, seed=9, sh_degree=1, bg=0.0, angle=1.1),
    "dense": dict(P=1500, W=48, H=48, logscale=-1.5, seed=4, sh_degree=2, bg=1.0),          # tiles with > 256 instances
    "deg0": dict(P=300, W=64, H=40, logscale=-2.0, seed=5, sh_degree=0, bg=0.5, scale_modifier=1.7),
    "precomp": dict(P=350, W=80, H=48, logscale=-2.3, seed=6, sh_degree=3, bg=1.0, precomp=True),
}


def make(name, kw):
    kw = dict(kw)
    precomp = kw.pop("precomp", False)
    smod = kw.pop("scale_modifier", 1.0)
    sc = util.scene(**kw)
    if precomp:
        # colours and covariances supplied by the caller (override_color / compute_cov3D_python paths)
        g = torch.Generator().manual_seed(17)
        sc["colors_precomp"] = torch.rand((sc["means3D"].shape[0], 3), generator=g)
        S = torch.diag_embed(sc["scales"])
        q = sc["rotations"]
        r, x, y, z = q.unbind(1)
        R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                         2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                         2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], dim=1).view(-1, 3, 3)
        Sig = R @ S @ S @ R.transpose(1, 2)
        sc["cov3D_precomp"] = torch.stack([Sig[:, 0, 0], Sig[:, 0, 1], Sig[:, 0, 2], Sig[:, 1, 1], Sig[:, 1, 2], Sig[:, 2, 2]], dim=1)
        sc["shs"] = None; sc["scales"] = None; sc["rotations"] = None
    refC = build_ref.load()
    scd = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in sc.items()}
    r = util.ref_forward(refC, scd, scale_modifier=smod)
    gen = torch.Generator().manual_seed(5)
    gc = torch.randn(r["color"].shape, generator=gen)
    runs = [util.ref_backward(refC, scd, r, gc.cuda(), scale_modifier=smod) for _ in range(3)]
    torch.cuda.synchronize()
    blob = {"meta/scale_modifier": np.float32(smod), "meta/sh_degree": np.int32(sc["sh_degree"]),
            "meta/W": np.int32(sc["W"]), "meta/H": np.int32(sc["H"]), "meta/tanfovx": np.float64(sc["tanfovx"]),
            "meta/tanfovy": np.float64(sc["tanfovy"]), "meta/num_rendered": np.int64(r["num_rendered"])}
    for k in ("bg", "means3D", "opacities", "scales", "rotations", "shs", "viewmatrix", "projmatrix", "campos",
              "colors_precomp", "cov3D_precomp"):
        if sc.get(k) is not None:
            blob["in/" + k] = sc[k].numpy()
    vis = (r["radii"] > 0).cpu()
    blob["out/radii"] = r["radii"].cpu().numpy()
    for k in ("depths", "means2D", "conic_opacity", "cov3D", "rgb", "tiles_touched"):
        v = r[k].cpu()
        m = vis if v.dim() == 1 else vis[:, None].expand_as(v)
        blob["out/" + k] = torch.where(m, v, torch.zeros_like(v)).numpy()     # invisible entries are uninitialised memory
    for k in ("keys", "point_list", "ranges", "n_contrib", "final_T", "color", "depth"):
        blob["out/" + k] = r[k].cpu().numpy()
    blob["in/dL_dpix"] = gc.numpy()
    for k in runs[0]:
        if runs[0][k].numel() == 0:
            continue
        a = torch.stack([x[k].double().cpu() for x in runs])
        blob["g/" + k] = a.mean(0).numpy()
        blob["gspread/" + k] = np.float64((a.max(0).values - a.min(0).values).abs().max() / (a.abs().max() + 1e-30))
    return blob


def main():
    out = os.path.join(ROOT, "gpurun_out", "golden")
    os.makedirs(out, exist_ok=True)
    for name, kw in CASES.items():
        blob = make(name, kw)
        np.savez_compressed(os.path.join(out, f"raster_{name}.npz"), **blob)
        print(name, "R =", int(blob["meta/num_rendered"]), "max inst/tile =",
              int((blob["out/ranges"][:, 1].astype(np.int64) - blob["out/ranges"][:, 0]).max()))


if __name__ == "__main__":
    main()
