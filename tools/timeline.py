"""GPU timeline of one training iteration (torch.profiler / CUPTI): busy time per kernel name, idle gaps.
usage (on a GPU box): python tools/timeline.py [--views 2]"""
import argparse, os, sys, collections
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from torch.profiler import profile, ProfilerActivity

ap = argparse.ArgumentParser(); ap.add_argument("--views", type=int, default=2); a = ap.parse_args()
dev = "cuda:0"; torch.cuda.set_device(0)
import fdgs
from fdgs import _lib as L
from fdgs.render import render as render_fn
from fdgs.dp import train_iteration
from fdgs.losses import l1_view_loss as fused_loss
L.load()
cfg, args, pc, net, cams, gts_host = bench.build_scene("dynerf", "fdgs", dev, a.views, 0)
bucket = bench.make_bucket(pc, net, dev)
bg = torch.ones(3, device=dev)
gts = [g.to(dev) for g in gts_host]
for _ in range(3):
    train_iteration(render_fn, pc, cams, gts, bg, bucket, 0, 1, loss_fn=fused_loss, speculative=True)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    train_iteration(render_fn, pc, cams, gts, bg, bucket, 0, 1, loss_fn=fused_loss, speculative=True)
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
evs.sort(key=lambda e: e.time_range.start)
t0, t1 = evs[0].time_range.start, max(e.time_range.end for e in evs)
busy = collections.defaultdict(float); cnt = collections.Counter()
end = t0; idle = 0.0; gaps = []
for e in evs:
    s, f = e.time_range.start, e.time_range.end
    if s > end:
        idle += s - end; gaps.append((s - end, prev, e.name[:50]))
    end = max(end, f); prev = e.name[:50]
    key = e.name[:60] if "at::native" not in e.name else "at:" + e.name[-150:]
    busy[key] += f - s; cnt[key] += 1
print(f"views {a.views}: span {t1 - t0:.0f} us, idle {idle:.0f} us ({100 * idle / (t1 - t0):.1f} %)")
for k, v in sorted(busy.items(), key=lambda x: -x[1])[:60]:
    print(f"{v:9.1f} us  x{cnt[k]:<4d} {k}")
print("largest gaps (us, after, before):")
for g in sorted(gaps, reverse=True)[:12]:
    print(f"  {g[0]:7.1f}  {g[1]}  ->  {g[2]}")
