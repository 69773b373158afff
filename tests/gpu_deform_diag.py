"""Diagnostic: fused deformation vs reference ops on GPU vs FP64 CPU arbiter at a BASELINE shape."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import fdgs  # noqa
from fdgs.configs import CONFIGS
import test_gpu_deform as T
import util, ref_port, copy

cfgname = sys.argv[1] if len(sys.argv) > 1 else "dnerf"
P = int(sys.argv[2]) if len(sys.argv) > 2 else 20011
c = CONFIGS[cfgname]
args = c["args"]()
ours, ref = T._pair(args)
aabb = ([c["extent"]] * 3, [-c["extent"]] * 3)
ours.deformation_net.set_aabb(*aabb); ref.deformation_net.set_aabb(*aabb)
ref64 = copy.deepcopy(ref).cpu().double()
g = torch.Generator().manual_seed(11)
inp = dict(xyz=(torch.rand((P, 3), generator=g) * 2 - 1) * c["extent"], scales=torch.randn((P, 3), generator=g),
           rot=torch.randn((P, 4), generator=g), opacity=torch.randn((P, 1), generator=g), shs=torch.randn((P, 16, 3), generator=g))
t = torch.full((P, 1), 0.4375)
with torch.no_grad():
    outs_r = ref64(*(inp[k].double() for k in ("xyz", "scales", "rot", "opacity", "shs")), t.double())
ws = [torch.randn(o.shape, generator=g) for o in outs_r]
o64, l64 = T._run(ref64, {k: v.double() for k, v in inp.items()}, t.double(), [w.double() for w in ws])
inpd = {k: v.cuda() for k, v in inp.items()}
o_r, l_r = T._run(ref, inpd, t.cuda(), [w.cuda() for w in ws])
o_o, l_o = T._run(ours, inpd, t.cuda(), [w.cuda() for w in ws])
for i in range(5):
    print(f"out[{i}] ours-vs-f64 {util.rel_err(o_o[i].detach(), o64[i].detach()):.3e}  refgpu-vs-f64 {util.rel_err(o_r[i].detach(), o64[i].detach()):.3e}")
ga, gb, g6 = l_o["xyz"].grad.cpu().double(), l_r["xyz"].grad.cpu().double(), l64["xyz"].grad
print(f"xyz grad ours-vs-f64 {util.rel_err(ga, g6):.3e}  refgpu-vs-f64 {util.rel_err(gb, g6):.3e}  ours-vs-refgpu {util.rel_err(ga, gb):.3e}")
d = (ga - g6).abs().max(dim=1).values
idx = torch.argsort(d, descending=True)[:6]
for i in idx.tolist():
    q = (inp["xyz"][i] - 1.3 * 0 - c["extent"]) * (2.0 / (-2 * c["extent"])) - 1.0
    print(i, "xyz", inp["xyz"][i].tolist(), "ours", ga[i].tolist(), "ref", gb[i].tolist(), "f64", g6[i].tolist())
    for l, m in enumerate(args.multires):
        W = 64 * m
        print("   level", l, "ix", (((q + 1) / 2) * (W - 1)).tolist())
dd = (gb - g6).abs().max(dim=1).values
print("count ours err > 1e-3:", int((d > 1e-3).sum()), " refgpu err > 1e-3:", int((dd > 1e-3).sum()), "median ours", float(d.median()), "median ref", float(dd.median()))
worst = 0
for (k, p), (_, q_), (_, r_) in zip(ours.named_parameters(), ref.named_parameters(), ref64.named_parameters()):
    if r_.grad is None: continue
    ea, eb = util.rel_err(p.grad, r_.grad), util.rel_err(q_.grad, r_.grad)
    if ea > 1e-5 or eb > 1e-5:
        print(f"param {k}: ours-vs-f64 {ea:.3e} refgpu-vs-f64 {eb:.3e}")
