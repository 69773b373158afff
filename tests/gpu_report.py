"""Diagnostic (not a pytest file): runs ours vs the reference CUDA build vs the CPU oracle on a few
scenes and prints every difference category.  Usage on the GPU box: python tests/gpu_report.py"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import util  # noqa: E402
import build_ref  # noqa: E402


def cmp_bits(name, a, b):
    a = a.detach().cpu().contiguous(); b = b.detach().cpu().contiguous()
    if a.shape != b.shape:
        print(f"   {name}: SHAPE {tuple(a.shape)} vs {tuple(b.shape)}")
        return
    if a.dtype.is_floating_point:
        ne = (a.view(torch.int32) != b.view(torch.int32)) if a.dtype == torch.float32 else (a != b)
    else:
        ne = a != b
    n = int(ne.sum())
    extra = ""
    if n and a.dtype.is_floating_point:
        extra = f" max|d|={float((a - b).abs().max()):.3e} rel={util.rel_err(a, b):.3e}"
    print(f"   {name}: {n} / {a.numel()} differ{extra}")


def report(title, sc, refC, with_oracle=True):
    print(f"== {title}: P={sc['means3D'].shape[0]} {sc['W']}x{sc['H']}")
    dev = "cuda"
    scd = {k: (v.to(dev) if isinstance(v, torch.Tensor) else v) for k, v in sc.items()}
    o = util.ours_forward(scd)
    torch.cuda.synchronize()
    print("   ours R =", o["num_rendered"])
    if refC is not None:
        r = util.ref_forward(refC, scd)
        torch.cuda.synchronize()
        print("   ref  R =", r["num_rendered"])
        vis = r["radii"] > 0
        cmp_bits("radii", o["radii"], r["radii"])
        for k in ("depths", "means2D", "conic_opacity", "rgb", "cov3D"):
            m = vis if o[k].dim() == 1 else vis[:, None].expand_as(o[k])
            cmp_bits(k + "[visible]", torch.where(m, o[k], torch.zeros_like(o[k])), torch.where(m, r[k], torch.zeros_like(r[k])))
        cmp_bits("tiles_touched", o["tiles_touched"], r["tiles_touched"])
        if o["num_rendered"] == r["num_rendered"]:
            cmp_bits("keys", o["keys"], r["keys"])
            cmp_bits("point_list", o["point_list"], r["point_list"])
        cmp_bits("ranges", o["ranges"], r["ranges"])
        cmp_bits("n_contrib", o["n_contrib"], r["n_contrib"])
        cmp_bits("final_T", o["final_T"], r["final_T"])
        cmp_bits("color", o["color"], r["color"])
        cmp_bits("depth", o["depth"], r["depth"])
        # backward
        gen = torch.Generator().manual_seed(5)
        gc = torch.randn(o["color"].shape, generator=gen).to(dev)
        _, _, _, go = util.ours_autograd(scd, gc)
        gr = util.ref_backward(refC, scd, r, gc)
        gr2 = util.ref_backward(refC, scd, r, gc)
        for k, rk in (("means3D", "means3D"), ("means2D", "means2D"), ("opacities", "opacities"), ("scales", "scales"),
                      ("rotations", "rotations"), ("shs", "sh")):
            if go.get(k) is None:
                continue
            a, b = go[k].reshape(-1), gr[rk].reshape(-1)
            spread = util.rel_err(gr2[rk], gr[rk])
            print(f"   grad {k}: rel err vs ref {util.rel_err(a, b):.3e}   (ref run-to-run {spread:.3e})")
    if with_oracle:
        t = time.time()
        c = util.oracle_forward(sc)
        print(f"   oracle f32 R = {c['num_rendered']} ({time.time() - t:.1f}s)")
        cmp_bits("radii vs oracle", o["radii"], torch.tensor(c["radii"]))
        cmp_bits("depths vs oracle", o["depths"] * (o["radii"] > 0), torch.tensor(c["depths"]).cuda() * (o["radii"] > 0))
        if o["num_rendered"] == c["num_rendered"]:
            cmp_bits("keys vs oracle", o["keys"], torch.tensor(c["keys"].astype(np.int64)))
            cmp_bits("point_list vs oracle", o["point_list"], torch.tensor(c["point_list"].astype(np.int32)))
        cmp_bits("ranges vs oracle", o["ranges"], torch.tensor(c["ranges"].astype(np.int32)))
        print("   color vs oracle rel:", util.rel_err(o["color"], c["color"]),
              " n_contrib diff:", int((o["n_contrib"].cpu() != torch.tensor(c["n_contrib"].astype(np.int32))).sum()))


def main():
    refC = build_ref.load() if build_ref.available() else None
    print("reference build:", "available" if refC is not None else "MISSING")
    report("tiny", util.scene(P=2000, W=200, H=120, logscale=-3.0), refC)
    report("ragged", util.scene(P=4096, W=333, H=217, logscale=-2.5, seed=9, bg=0.0), refC)
    report("big splats", util.scene(P=500, W=256, H=256, logscale=-0.5, seed=4), refC)
    report("dnerf50k", util.scene(P=50000, W=800, H=800, logscale=-4.605, seed=0), refC, with_oracle=False)
    report("dynerf300k", util.scene(P=300000, W=1352, H=1014, fovx=1.2, extent=2.0, logscale=-4.828, seed=0), refC,
           with_oracle=False)


if __name__ == "__main__":
    main()
