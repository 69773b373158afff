"""bench.py — fwd+bwd train iters/s (and render FPS) of the 4D-Gaussian hot path on N B200s.

One "step" (default `--mode train`) = one training iteration of a BASELINE.json configuration (default: config 4,
DyNeRF shape): for each of the B views of the batch  deformation fwd -> exp/normalize/sigmoid -> rasterize fwd ->
L1 vs a fixed random GT -> backward through all of it,  then (N>1) one NCCL all-reduce of the flat gradient bucket.
Views are sharded view i -> rank i mod N (strong scaling: the batch is fixed).  Optimizer step, densification, data
loading and logging are outside the measured unit; the channel-last plane copy the deformation keeps
across views is invalidated once per step, as a real optimizer step would.

`--mode render-sweep` (BASELINE.json config 5, modelled on the reference's render.py:56-70) renders `--frames` (300)
timestamps t = k/frames forward-only under no_grad and reports frames/s, with and without the per-frame device->host
copy of the image that render.py's `to8b(rendering)` performs.  Frames are sharded over the ranks (replicas only).

  python bench.py [--gpus N --steps K --warmup W] [--impl fdgs|reference] [--config dynerf] [--mode train|render-sweep]

`--impl reference` runs the reference arm: the reference's own deformation module (scene/deformation.py +
scene/hexplane.py, unmodified, loaded from baseline/_ref — or oracle/ref_port.py, its op-for-op port, when that copy is
absent; `config.reference_arm` says which) + the UNMODIFIED reference CUDA rasterizer compiled from its own sources
(oracle/_ref), i.e. what the reference gives on this GPU; none of the fdgs kernels is on that path.
"""
import argparse
import json
import math
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="fdgs", choices=["fdgs", "reference"])
    ap.add_argument("--config", default="dynerf")
    ap.add_argument("--mode", default="train", choices=["train", "render-sweep"])
    ap.add_argument("--frames", type=int, default=300, help="render-sweep: timestamps per step")
    ap.add_argument("--views", type=int, default=0, help="views per batch (default: the config's)")
    ap.add_argument("--points", type=int, default=0, help="Gaussians (default: the config's)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the render-FPS, phase-split and per-kernel passes")
    return ap.parse_args()


METRIC = {
    "dynerf": "fwd+bwd train iters/s (DyNeRF-shape, 300k G, batch 8 views)",
    "dnerf": "fwd+bwd train iters/s (D-NeRF-shape, 200k G, 800x800, batch 1 view)",
    "dnerf50k": "fwd+bwd train iters/s (D-NeRF bouncingballs-shape, 50k G, 800x800, batch 1 view)",
    "hypernerf": "fwd+bwd train iters/s (HyperNeRF-shape, 300k G, 536x960, batch 2 views)",
    "infer1080p": "fwd+bwd train iters/s (1M G, 1920x1080)",
}


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """nvidia-smi sampled every 200 ms DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "200"], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 7:
                continue
            try:
                sm.append(float(c[0])); mx.append(float(c[1]))
            except ValueError:
                continue
            for n, v in zip(names, c[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        self.f.close()
        try:
            os.unlink(self.f.name)
        except OSError:
            pass
        if sm:
            sm.sort()
            out.update(sm_mhz=sm[len(sm) // 2], sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        return out


# ------------------------------------------------------------------------------------------------ scene
def reference_net_factory():
    """(make_net, label): the reference's own deform_network when its files are installed, else the op-for-op port."""
    import ref_deform
    if ref_deform.available():
        return ref_deform.make_network, "scene/deformation.py + scene/hexplane.py (unmodified reference module, " + \
            ("/root/reference" if ref_deform.root().startswith("/root/reference") else "baseline/_ref") + ")"
    import ref_port
    return ref_port.ref_deform_network, "oracle/ref_port.py (op-for-op port; baseline/_ref not installed)"


def build_scene(cfg_name, impl, dev, views, points, make_gts=True, n_cams=None, time_den=None):
    import fdgs  # noqa: F401
    from fdgs import synth
    from fdgs.configs import CONFIGS
    cfg = dict(CONFIGS[cfg_name])
    if views:
        cfg["views"] = views
    if points:
        cfg["P"] = points
    args = cfg["args"]()
    if impl == "fdgs":
        from fdgs.deformation import deform_network as make_net
        cfg["net_impl"] = "fdgs.deformation.deform_network"
    else:
        make_net, cfg["net_impl"] = reference_net_factory()
    import ref_deform
    with ref_deform.quiet_stdout():      # the reference module prints from its constructor / set_aabb
        net = synth.seeded_deformation(make_net, args, seed=1234, device=dev)
        g = synth.make_gaussians(cfg["P"], extent=cfg["extent"], logscale=cfg["logscale"], seed=0, device=dev)
        pc = synth.SynthGaussians(g, net).to(dev)
        net.deformation_net.set_aabb(pc._xyz.detach().max(0).values.tolist(), pc._xyz.detach().min(0).values.tolist())
        net.to(dev)      # the reference's set_aabb re-creates the aabb Parameter on the CPU (scene/hexplane.py:118-122)
    B = cfg["views"]
    ncam = n_cams if n_cams else B
    cams = synth.make_cameras(cfg["W"], cfg["H"], cfg["fovx"], ncam, device=dev, time_den=time_den)
    gts_host = []
    if make_gts:
        gen = torch.Generator().manual_seed(99)
        gts_host = [torch.rand((3, cfg["H"], cfg["W"]), generator=gen).pin_memory() if dev != "cpu" else
                    torch.rand((3, cfg["H"], cfg["W"]), generator=gen) for _ in range(B)]
    return cfg, args, pc, net, cams, gts_host


def host_cameras(cams):
    """Pinned host copies of the per-view camera tensors (the e2e path uploads them every step,
    as gaussian_renderer/__init__.py:45-48 does)."""
    out = []
    for c in cams:
        out.append({k: getattr(c, k).cpu().pin_memory() for k in ("world_view_transform", "full_proj_transform", "camera_center")})
    return out


def upload_camera(cam, host, dev):
    from types import SimpleNamespace
    return SimpleNamespace(image_width=cam.image_width, image_height=cam.image_height, FoVx=cam.FoVx, FoVy=cam.FoVy,
                           time=cam.time, **{k: v.to(dev, non_blocking=True) for k, v in host.items()})


def make_bucket(pc, net, dev):
    from fdgs.dp import GradBucket
    params = [pc._xyz, pc._scaling, pc._rotation, pc._opacity, pc._features_dc, pc._features_rest]
    dn = net.deformation_net
    hm_names = ("pos_deform", "scales_deform", "rotations_deform", "opacity_deform", "shs_deform")
    a = dn.args
    active = [not a.no_dx, not a.no_ds, not a.no_dr, not a.no_do, not a.no_dshs]
    params += [p for planes in dn.grid.grids for p in planes]
    params += list(dn.feature_out.parameters())
    for name, on in zip(hm_names, active):
        if on:
            params += list(getattr(dn, name).parameters())
    bucket = GradBucket(params, extra={"viewspace": (pc._xyz.shape[0], 3)}, device=dev)
    from fdgs.dp import grads_in_place
    grads_in_place(net)      # fdgs deformation backward writes the bucket slots directly (no-op for the reference arm)
    return bucket


# ------------------------------------------------------------------------------------------------ roofline
SM_COUNT = 148
# Minimum warp instructions the blend inner loops need per pixel x Gaussian pair that CONTRIBUTES (derivation: DESIGN.md
# §3 "blend ceilings"): forward 25 (power 8, exp 2, alpha/threshold tests 5, T update 3, 4 accumulate FMAs,
# bookkeeping 3), backward 62 (forward recompute 14, T / accumulated colour recurrences 16, dL/dalpha 8, nine partials 19,
# its share of the cross-lane reduction 5).
BLEND_MIN_INSTR = {"k_blend_fwd": 25.0, "k_blend_bwd": 62.0}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(hbm=float(d["hbm_gbs"]), tensor=float(d.get("bf16_tflops_sustained", d["bf16_tflops"])),
                    sm_mhz=float(d.get("sm_max_mhz", 1965.0)), src="measured (MEASURED_PEAKS.json)")
    return dict(hbm=6650.0, tensor=1400.0, sm_mhz=1965.0, src="fallback (B200_PROFILING.md)")


def algorithmic_per_launch(kernel, s):
    """(bound, amount) of ALGORITHMIC traffic / work of one launch of `kernel` for one view of the workload
    s = dict(P, M, R, HW, T, F, Wn, heads, sum_out, hits).  Figures are stated in DESIGN.md §3."""
    P, M, R, HW, T = s["P"], s["M"], s["R"], s["HW"], s["T"]
    mlp_fwd_flop = 2.0 * P * (s["F"] * s["Wn"] + s["heads"] * s["Wn"] ** 2 + s["Wn"] * s["sum_out"])
    table = {
        "k_preprocess": ("hbm", P * (44 + 12 * M + 75) + 0.0),
        "k_tile_scan": ("hbm", T * 20.0),
        "k_scatter": ("hbm", P * 16.0 + R * 8.0),
        "k_tile_sort": ("hbm", R * (8 + 36 + 48) + 0.0),
        "k_blend_fwd": ("fp32_issue", float(s.get("hits") or 0)),
        "k_blend_bwd": ("fp32_issue", float(s.get("hits") or 0)),
        "k_preprocess_bwd": ("hbm", P * (301.0 - 192 + 12 * M) + P * (64.0 + 12 * M)),
        "k_deform_fwd": ("tensor", mlp_fwd_flop),
        "k_deform_fwd_tc": ("tensor", mlp_fwd_flop),
        "k_deform_bwd": ("tensor", 2.0 * mlp_fwd_flop),
        # the MLP backward is 2x the forward FLOPs: one forward's worth of data gradients, one of weight gradients
        # (the activation recompute inside k_deform_bwd_data is NOT counted as algorithmic work)
        "k_deform_bwd_data": ("tensor", mlp_fwd_flop),
        "k_deform_bwd_data_saved": ("tensor", mlp_fwd_flop),
        "k_deform_bwd_weights": ("tensor", mlp_fwd_flop),
        # sampler backward: dL/dfeat row + xyz in, d_xyz out, and the float4 reductions it sends to the L2-resident plane
        # gradients (6 planes: 3 x 4 + 3 x 2 taps of 16 B per 4-channel quad) counted as bytes moved
        "k_deform_sample_bwd": ("hbm", P * (4.0 * s["F"] + 24.0) + P * (s["F"] / 4.0) * 18 * 16.0),
    }
    return table.get(kernel)


HBM_BYTES_BLEND = {   # compulsory HBM bytes of the two blend kernels (second field of their roofline entries)
    "k_blend_fwd": lambda s: s["R"] * 48.0 + s["HW"] * 24.0 + s["T"] * 8.0,
    "k_blend_bwd": lambda s: s["R"] * 48.0 + s["HW"] * 20.0 + s["T"] * 8.0 + s["P"] * 44.0,
}

# dram__bytes_read.sum + dram__bytes_write.sum per launch from the `ncu --set full` capture of the DyNeRF workload
# (profiles/r02_ncu_traffic.json when committed, else the r01 capture below; P = 300k view).  Only for that workload.
NCU_TRAFFIC_BYTES = {
    "k_deform_fwd_tc": 0.075682e9 + 1.007662e9, "k_deform_bwd_data_saved": 0.100450e9 + 0.905332e9,
    "k_deform_bwd_weights": 1.987011e9 + 0.004113e9, "k_deform_sample_bwd": 0.049985e9 + 0.000285e9,
    "k_blend_bwd": 0.119234e9 + 0.003775e9, "k_blend_fwd": 0.089780e9 + 0.011699e9,
    "k_tile_sort": 0.025838e9 + 0.035497e9, "k_preprocess": 0.066367e9 + 0.003340e9, "k_preprocess_bwd": 0.093701e9 + 0.042599e9,
}
NCU_TRAFFIC_SOURCE = ["profiles/r01_ncu_full_summary.csv"]


def load_ncu_traffic():
    """Overrides the table above with profiles/r02_ncu_traffic.json ({kernel: bytes per launch}) when it is committed."""
    p = os.path.join(ROOT, "profiles", "r02_ncu_traffic.json")
    if os.path.exists(p):
        try:
            with open(p) as f:
                d = json.load(f)
            NCU_TRAFFIC_BYTES.clear()
            NCU_TRAFFIC_BYTES.update({k: float(v) for k, v in d.get("bytes_per_launch", {}).items()})
            NCU_TRAFFIC_SOURCE[0] = d.get("source", "profiles/r02_ncu_traffic.json")
        except Exception:
            pass


def roofline_entry(kernel, n, ms, shape, peaks):
    alg = algorithmic_per_launch(kernel, shape)
    if alg is None or n <= 0:
        return None
    bound, amount = alg
    sec = (ms / n) * 1e-3
    e = {"kernel": kernel, "bound": bound, "avg_launch_us": round(sec * 1e6, 2), "algorithmic_per_launch": amount}
    if bound == "hbm":
        ach, unit, peak = amount / sec / 1e9, "GB/s", peaks["hbm"]
    elif bound == "tensor":
        ach, unit, peak = amount / sec / 1e12, "TFLOP/s", peaks["tensor"]
        e["note"] = ("algorithmic FP32 FLOPs against the measured bf16 peak; the GEMMs run as 3xTF32 (3 passes at half the bf16 "
                     "rate), so the ceiling of frac for this precision scheme is 1/6 = 0.167")
        e["frac_of_3xtf32_ceiling"] = round(ach / peak * 6.0, 4)
    else:
        # FP32-issue bound: contributing pixel x Gaussian pairs per second against
        # SMs x 4 schedulers x 32 lanes x clock / (minimum warp instructions per contributing pair)
        ach, unit = amount / sec / 1e9, "Gpair/s"
        peak = SM_COUNT * 4 * 32 * peaks["sm_mhz"] * 1e6 / BLEND_MIN_INSTR[kernel] / 1e9
        e["min_instr_per_pair"] = BLEND_MIN_INSTR[kernel]
        e["pairs_tested_per_launch"] = shape.get("walked")
        hb = HBM_BYTES_BLEND[kernel](shape)
        e["hbm"] = {"achieved": round(hb / sec / 1e9, 2), "peak": peaks["hbm"], "unit": "GB/s", "frac": round(hb / sec / 1e9 / peaks["hbm"], 5)}
    e.update(achieved=round(ach, 3), peak=round(peak, 2), unit=unit, frac=round(ach / peak, 5), peak_source=peaks["src"])
    tr = NCU_TRAFFIC_BYTES.get(kernel) if (shape["P"] == 300000 and shape["Wn"] == 128) else None
    e["traffic"] = tr
    if tr is not None:
        e["traffic_source"] = NCU_TRAFFIC_SOURCE[0]
    return e


def rooflines_for(prof, shape, peaks, min_share=0.05):
    """prof: {kernel: (launches, total_ms)} from the in-library CUDA-event profiler.  Returns (entry of the dominant
    kernel, entries of every kernel above `min_share` of the library time, per-kernel table)."""
    if not prof:
        return None, [], {}
    tot = sum(v[1] for v in prof.values()) or 1.0
    per_kernel = {k: {"launches": v[0], "avg_us": round(1e3 * v[1] / max(v[0], 1), 2), "share": round(v[1] / tot, 4)}
                  for k, v in sorted(prof.items(), key=lambda kv: -kv[1][1])}
    entries = []
    for k, v in sorted(prof.items(), key=lambda kv: -kv[1][1]):
        if v[1] / tot < min_share:
            continue
        e = roofline_entry(k, v[0], v[1], shape, peaks)
        if e is not None:
            e["share"] = round(v[1] / tot, 4)
            entries.append(e)
    return (entries[0] if entries else None), entries, per_kernel


# ------------------------------------------------------------------------------------------------ CPU baseline
def cpu_baseline(cfg, args, n_views):
    """Bounded sample on the host cores: ONE view of the batch — the reference's PyTorch deformation (the real module
    when installed, else its port; all cores) fwd+bwd plus the CPU rasterizer restatement (oracle/raster_oracle.cpp,
    FP64 instantiation, 1 core) fwd+bwd — extrapolated to the B-view iteration."""
    import raster_oracle
    import ref_deform
    from fdgs import synth
    ncores = os.cpu_count() or 1
    torch.set_num_threads(ncores)
    P = cfg["P"]
    make_net, label = reference_net_factory()
    with ref_deform.quiet_stdout():
        net = synth.seeded_deformation(make_net, args, seed=1234, device="cpu")
        g = synth.make_gaussians(P, extent=cfg["extent"], logscale=cfg["logscale"], seed=0, device="cpu")
        net.deformation_net.set_aabb(g["xyz"].max(0).values.tolist(), g["xyz"].min(0).values.tolist())
    leaves = {k: v.clone().requires_grad_(True) for k, v in g.items()}
    shs = torch.cat([leaves["f_dc"], leaves["f_rest"]], dim=1)
    t = torch.full((P, 1), 0.25)
    t0 = time.perf_counter()
    outs = net(leaves["xyz"], leaves["scaling"], leaves["rotation"], leaves["opacity"], shs, t)
    sum(o.sum() for o in outs).backward()
    t_def = time.perf_counter() - t0
    cam = synth.make_camera(cfg["W"], cfg["H"], cfg["fovx"], angle=0.0, time=0.25)
    xyz, sc, rot, op, sh = [o.detach() for o in outs]
    gen = torch.Generator().manual_seed(5)
    dpix = torch.randn((3, cfg["H"], cfg["W"]), generator=gen)
    t0 = time.perf_counter()
    raster_oracle.rasterize(W=cfg["W"], H=cfg["H"], tanfovx=math.tan(cam.FoVx * 0.5), tanfovy=math.tan(cam.FoVy * 0.5),
                            bg=torch.ones(3).numpy(), means3D=xyz.numpy(), opacities=torch.sigmoid(op).numpy(),
                            viewmatrix=cam.world_view_transform.numpy(), projmatrix=cam.full_proj_transform.numpy(),
                            campos=cam.camera_center.numpy(), shs=sh.numpy(), scales=torch.exp(sc).numpy(),
                            rotations=torch.nn.functional.normalize(rot).numpy(), sh_degree=3, dL_dpix=dpix.numpy(), f64=True)
    t_ras = time.perf_counter() - t0
    per_iter = (t_def + t_ras) * n_views
    return {"value": round(1.0 / per_iter, 6), "unit": "iters/s", "cores": ncores, "kind": "port",
            "sample": f"1 of {n_views} views at full size (P={P}, {cfg['W']}x{cfg['H']}): {label} fwd+bwd "
                      f"{t_def:.2f}s on {ncores} threads + C++ raster restatement fwd+bwd {t_ras:.2f}s on 1 thread; x{n_views}",
            "deform_s_per_view": round(t_def, 3), "raster_s_per_view": round(t_ras, 3)}


# ------------------------------------------------------------------------------------------------ phase split
def phase_split(impl, pc, net, cam, bg, refC=None, reps=3):
    """Per-view device time (CUDA events, ms) of the two operators in isolation, same method for both arms:
    deformation fwd / bwd (the module called as render() calls it) and rasterizer fwd / bwd (the autograd Function on
    detached activated inputs).  The reference arm calls the reference module and the unmodified reference rasterizer."""
    import ref_port
    from fdgs.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    dev = pc._xyz.device
    P = pc._xyz.shape[0]
    ev = lambda: torch.cuda.Event(enable_timing=True)
    t_arg = cam.time if impl == "fdgs" else torch.tensor(cam.time).to(dev).repeat(P, 1)
    settings = GaussianRasterizationSettings(
        image_height=int(cam.image_height), image_width=int(cam.image_width), tanfovx=math.tan(cam.FoVx * 0.5),
        tanfovy=math.tan(cam.FoVy * 0.5), bg=bg, scale_modifier=1.0, viewmatrix=cam.world_view_transform,
        projmatrix=cam.full_proj_transform, sh_degree=pc.active_sh_degree, campos=cam.camera_center, prefiltered=False, debug=False)
    rast = GaussianRasterizer(settings) if impl == "fdgs" else ref_port.RefGaussianRasterizer(refC, settings)
    acc = dict(deform_fwd_ms=0.0, deform_bwd_ms=0.0, raster_fwd_ms=0.0, raster_bwd_ms=0.0)
    gen = torch.Generator(device=dev).manual_seed(7)
    for r in range(reps + 1):
        feats = pc.get_features
        e = [ev() for _ in range(4)]
        e[0].record()
        outs = net(pc._xyz, pc._scaling, pc._rotation, pc._opacity, feats, t_arg)
        e[1].record()
        gs = [torch.randn(o.shape, generator=gen, device=dev) for o in outs]
        e[2].record()
        torch.autograd.backward(outs, gs)
        e[3].record()
        with torch.no_grad():
            xyz, sc, rot, op, sh = [o.detach() for o in outs]
            sc, rot, op = torch.exp(sc), torch.nn.functional.normalize(rot), torch.sigmoid(op)
        leaves = [t.clone().requires_grad_(True) for t in (xyz, op, sh, sc, rot)]
        m2d = torch.zeros_like(xyz, requires_grad=True)
        gimg = torch.randn((3, int(cam.image_height), int(cam.image_width)), generator=gen, device=dev)
        e4, e5, e6 = ev(), ev(), ev()
        e4.record()
        color, radii, depth = rast(means3D=leaves[0], means2D=m2d, opacities=leaves[1], shs=leaves[2], scales=leaves[3],
                                   rotations=leaves[4])
        e5.record()
        color.backward(gimg)
        e6.record()
        torch.cuda.synchronize()
        if r == 0:
            continue        # warm-up
        acc["deform_fwd_ms"] += e[0].elapsed_time(e[1]); acc["deform_bwd_ms"] += e[2].elapsed_time(e[3])
        acc["raster_fwd_ms"] += e4.elapsed_time(e5); acc["raster_bwd_ms"] += e5.elapsed_time(e6)
    out = {k: round(v / reps, 4) for k, v in acc.items()}
    out["deform_ms"] = round(out["deform_fwd_ms"] + out["deform_bwd_ms"], 4)
    out["raster_ms"] = round(out["raster_fwd_ms"] + out["raster_bwd_ms"], 4)
    out["how"] = "per view, operators timed in isolation with CUDA events (mean of %d after 1 warm-up)" % reps
    return out


# ------------------------------------------------------------------------------------------------ main
def main():
    a = parse()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if a.impl == "reference" and rank != 0:
        return 0          # the reference has no multi-GPU path: rank 0 alone runs the whole batch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the fdgs hot path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = f"cuda:{local_rank}"
    use_dist = world > 1 and a.impl == "fdgs"
    if use_dist:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL writes its version banner to file descriptor 1: keep stdout to the ONE JSON line
        sys.stdout.flush()
        json_fd = os.dup(1)
        os.dup2(2, 1)
        sys.stdout = os.fdopen(json_fd, "w")
        dist.init_process_group("nccl", device_id=torch.device(dev))
    eff_world, eff_rank = (world, rank) if use_dist else (1, 0)

    refC = None
    if a.impl == "reference":
        import build_ref
        if not build_ref.available():
            print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref (reference rasterizer built from its own sources) is missing"}))
            return 0
        refC = build_ref.load()
        import ref_port
        from fdgs.diff_gaussian_rasterization import GaussianRasterizationSettings as SettingsTuple   # plain NamedTuple

        def render_fn(cam, pc, bg_color=None, stage="fine"):
            return ref_port.ref_render(refC, SettingsTuple, cam, pc, bg_color, stage=stage)
        launch_count = lambda: 0
        L = None
        fused_loss = None
    else:
        import fdgs  # noqa: F401
        from fdgs import _lib as L
        from fdgs.render import render as render_fn
        L.load()
        launch_count = L.launch_count
        from fdgs.losses import l1_view_loss as fused_loss     # the reference arm keeps the ATen loss

    if a.mode == "render-sweep":
        return render_sweep(a, dev, render_fn, launch_count, use_dist, eff_rank, eff_world, world)

    cfg, args, pc, net, cams, gts_host = build_scene(a.config, a.impl, dev, a.views, a.points)
    from fdgs.dp import train_iteration, shard_views
    B = cfg["views"]
    bucket = make_bucket(pc, net, dev)
    bg = torch.ones(3, device=dev)
    gts_dev = [g.to(dev) for g in gts_host]
    cams_host = host_cameras(cams)
    mine = shard_views(B, eff_rank, eff_world)

    def barrier():
        if use_dist:
            dist.barrier()
        torch.cuda.synchronize()

    overlap_sh = os.environ.get("FDGS_OVERLAP_SH", "0") != "0"    # A/B switch for the early SH all-reduce (N > 1)
    # fdgs arm: one-shot rasterizer forward (no per-view host sync on num_rendered; FDGS_ONE_SHOT=0 = the exact path)
    one_shot = a.impl == "fdgs" and os.environ.get("FDGS_ONE_SHOT", "1") != "0"

    def optimizer_stepped():
        # what an optimizer step between two iterations implies for the fdgs arm: the cached channel-last plane copy is
        # stale and k_pack_plane runs again in the step's first deformation forward (the bench never steps the optimizer)
        dn = net.deformation_net
        if hasattr(dn, "_packed_key"):
            dn._packed_key = None

    def step_resident():
        optimizer_stepped()
        train_iteration(render_fn, pc, cams, gts_dev, bg, bucket, eff_rank, eff_world, loss_fn=fused_loss, overlap_sh=overlap_sh,
                        speculative=one_shot)

    from fdgs.dp import ViewStager
    stager = ViewStager([gts_host[i] if i in mine else None for i in range(B)], dev)

    class StepCameras:
        """This step's cameras: view i's matrices are uploaded (pinned -> device) when train_iteration reaches view i."""

        def __len__(self):
            return len(cams)

        def __getitem__(self, i):
            return upload_camera(cams[i], cams_host[i], dev)

    def step_e2e():
        # host -> device: this step's GT images (pinned -> device on the stager's copy stream, one event per view, so the
        # copy of view i+1 runs under the kernels of view i) and camera tensors; device -> host: the loss
        optimizer_stepped()
        stager.upload(gts_host, mine)
        loss, _, _ = train_iteration(render_fn, pc, StepCameras(), stager, bg, bucket, eff_rank, eff_world, read_loss=True,
                                     loss_fn=fused_loss, overlap_sh=overlap_sh, speculative=one_shot)
        assert stager.nbytes == sum(gts_host[i].numel() * 4 for i in mine)
        return loss

    def timed(fn, steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if use_dist:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms)

    for _ in range(max(a.warmup, 3)):
        step_resident()
    sampler = ClockSampler(local_rank) if eff_rank == 0 else None
    n0 = launch_count()
    ms = timed(step_resident, a.steps)
    launches = launch_count() - n0
    clocks = sampler.stop() if sampler else {}
    if one_shot:
        # the timed steps never read anything back: examine their counts now (a step that overflowed its instance
        # workspace rendered nothing and would make the number above invalid)
        from fdgs.diff_gaussian_rasterization import check_pending
        if not check_pending(block=True):
            raise SystemExit("bench: a one-shot render overflowed its workspace inside the timed region; rerun with FDGS_ONE_SHOT=0")
    for _ in range(2):
        step_e2e()
    ms_e2e = timed(step_e2e, a.steps)

    extras = {}
    roof, roofs, per_kernel = None, [], {}
    if not a.no_extras:
        # render FPS: forward only, no autograd (render.py:56-70 of the reference)
        with torch.no_grad():
            for i in mine[:1]:
                render_fn(cams[i], pc, bg_color=bg, stage="fine")
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier(); e0.record()
            for _ in range(3):
                for i in mine:
                    render_fn(cams[i], pc, bg_color=bg, stage="fine")
            e1.record(); barrier()
            fps_ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
            if use_dist:
                dist.all_reduce(fps_ms, op=dist.ReduceOp.MAX)
            extras["render_fps"] = round(3 * B / (float(fps_ms) * 1e-3), 2)
        if eff_rank == 0 and mine:
            try:
                extras["phase_split"] = phase_split(a.impl, pc, net, cams[mine[0]], bg, refC=refC)
            except Exception as ex:
                extras["phase_split"] = {"error": repr(ex)}
        if L is not None:
            # per-kernel CUDA-event timing inside the library, on the launching stream.  Every rank runs the two
            # iterations (they contain the gradient all-reduce, a collective); rank 0 alone records and reports.
            if eff_rank == 0:
                L.profile_collect()
                L.profile_enable(True)
            for _ in range(2):
                step_resident()
        if L is not None and eff_rank == 0:
            prof = L.profile_collect()
            L.profile_enable(False)
            P = cfg["P"]
            a_ = args
            heads_on = [not a_.no_dx, not a_.no_ds, not a_.no_dr, not a_.no_do, not a_.no_dshs]
            outs = (3, 3, 4, 1, 48)
            from fdgs.diff_gaussian_rasterization import count_pairs
            pairs = count_pairs(render_fn, cams[mine[0]], pc, bg)
            shape = dict(P=P, M=16, R=pairs["num_rendered"], HW=cfg["W"] * cfg["H"],
                         T=((cfg["W"] + 15) // 16) * ((cfg["H"] + 15) // 16),
                         F=a_.kplanes_config["output_coordinate_dim"] * len(a_.multires), Wn=a_.net_width,
                         heads=sum(heads_on), sum_out=sum(o for o, on in zip(outs, heads_on) if on),
                         hits=pairs["contributing"], walked=pairs["tested"])
            load_ncu_traffic()
            roof, roofs, per_kernel = rooflines_for(prof, shape, load_peaks())
            extras["num_rendered_per_view"] = pairs["num_rendered"]
            extras["blend_pairs_per_view"] = {"tested": pairs["tested"], "contributing": pairs["contributing"],
                                              "upper_bound_256R": 256 * pairs["num_rendered"]}

    if eff_rank == 0:
        P = cfg["P"]
        h2d = sum(gts_host[i].numel() * 4 + 35 * 4 for i in mine)
        metric = METRIC.get(a.config, f"fwd+bwd train iters/s ({a.config}-shape)")
        if a.views or a.points:
            metric = f"fwd+bwd train iters/s ({a.config}-shape, {P} G, batch {B} views)"
        line = {
            "metric": metric, "value": round(a.steps / (ms * 1e-3), 4),
            "unit": "iters/s", "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
            "ms_per_step": round(ms / a.steps, 3), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{a.config}: {P} Gaussians, {cfg['W']}x{cfg['H']}, batch {B} views, deformation+rasterizer fwd+bwd"
                                   + (", NCCL grad all-reduce" if use_dist else ""),
                       "views_per_rank": len(mine), "parallelism": f"view-dp{world}",
                       "l2": "working set per step >> L2 (>=400 MB per view streamed; inputs larger than L2)",
                       "grad_bucket_bytes": bucket.nbytes(), "deformation": cfg["net_impl"],
                       "plane_cache": "invalidated once per step (as after an optimizer step)",
                       "rasterizer_forward": ("one-shot (no host sync on num_rendered; counts verified after the timed region: "
                                              "no overflow)" if one_shot else "plan + render with the reference's blocking read of num_rendered")},
            "clocks": clocks,
            "e2e": {"value": round(a.steps / (ms_e2e * 1e-3), 4), "unit": "iters/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": 4},
            "gpu_launches": int(launches),
        }
        line.update(extras)
        if a.impl == "reference":
            line["impl"] = "reference"
            line["gpu_launches"] = 0
            line["config"]["reference_arm"] = (cfg["net_impl"] + " on the GPU + unmodified reference CUDA rasterizer "
                                               "(oracle/_ref, sm_100); single GPU runs the whole batch")
            line["e2e"] = {"value": line["e2e"]["value"], "unit": "iters/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4}
        else:
            line["roofline"] = roof
            line["rooflines"] = roofs
            line["kernels"] = per_kernel
        if world == 1 and not a.no_cpu_baseline:
            try:
                line["cpu_baseline"] = cpu_baseline(cfg, args, B)
            except Exception as ex:   # the CPU leg must never take the GPU numbers down with it
                line["cpu_baseline"] = {"error": repr(ex)}
        print(json.dumps(line), flush=True)
    if use_dist:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def render_sweep(a, dev, render_fn, launch_count, use_dist, eff_rank, eff_world, world):
    """BASELINE.json config 5: `--frames` timestamps t = k/frames of one scene, deformation + rasterize forward only.
    One step = the whole sweep; frames are sharded k -> rank k mod N (no collective on the data path)."""
    cfg_name = a.config if a.config != "dynerf" else "infer1080p"
    frames = a.frames
    cfg, args, pc, net, cams, _ = build_scene(cfg_name, a.impl, dev, a.views, a.points, make_gts=False, n_cams=frames,
                                              time_den=frames)
    bg = torch.ones(3, device=dev)
    mine = [k for k in range(frames) if k % eff_world == eff_rank]
    host_img = torch.empty((3, cfg["H"], cfg["W"]), dtype=torch.float32).pin_memory()

    def barrier():
        if use_dist:
            dist.barrier()
        torch.cuda.synchronize()

    def sweep(copy_back):
        with torch.no_grad():
            for k in mine:
                img = render_fn(cams[k], pc, bg_color=bg, stage="fine")["render"]
                if copy_back:        # render.py:60 `to8b(rendering)`: a blocking device -> host copy of every frame
                    host_img.copy_(img)

    def timed(fn, steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier(); e0.record()
        for _ in range(steps):
            fn()
        e1.record(); barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if use_dist:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms)

    with torch.no_grad():
        for k in mine[:max(a.warmup, 3)]:
            render_fn(cams[k], pc, bg_color=bg, stage="fine")
    sampler = ClockSampler(int(dev.split(":")[1])) if eff_rank == 0 else None
    n0 = launch_count()
    ms = timed(lambda: sweep(False), a.steps)
    launches = launch_count() - n0
    clocks = sampler.stop() if sampler else {}
    ms_copy = timed(lambda: sweep(True), a.steps)
    if eff_rank == 0:
        P = cfg["P"]
        fps = frames * a.steps / (ms * 1e-3)
        fps_copy = frames * a.steps / (ms_copy * 1e-3)
        img_bytes = 3 * cfg["H"] * cfg["W"] * 4
        line = {
            "metric": f"render FPS ({P // 1000}k-Gaussian {cfg['W']}x{cfg['H']}, {frames}-timestamp sweep, deformation+rasterize forward)",
            "value": round(fps, 3), "unit": "frames/s", "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
            "ms_per_step": round(ms / a.steps, 3), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{cfg_name}: {P} Gaussians, {cfg['W']}x{cfg['H']}, {frames} timestamps t=k/{frames}, forward only (no autograd)",
                       "frames_per_rank": len(mine), "parallelism": f"frame-replicas{world}", "deformation": cfg["net_impl"],
                       "l2": "per-frame working set >> L2 (1M Gaussians x 236 B inputs + instance lists)"},
            "clocks": clocks,
            "e2e": {"value": round(fps_copy, 3), "unit": "frames/s", "h2d_bytes_per_step": frames * 35 * 4 // max(world, 1),
                    "d2h_bytes_per_step": len(mine) * img_bytes,
                    "note": "every frame copied device -> pinned host, blocking, as render.py:60 `to8b(rendering)` does"},
            "gpu_launches": int(launches),
        }
        if a.impl == "reference":
            line["impl"] = "reference"
            line["gpu_launches"] = 0
            line["config"]["reference_arm"] = cfg["net_impl"] + " on the GPU + unmodified reference CUDA rasterizer (oracle/_ref, sm_100)"
        print(json.dumps(line), flush=True)
    if use_dist:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
