"""GPU timeline of one view-parallel training iteration on rank 0 (torch.profiler / CUPTI) under torchrun:
kernel busy time, the NCCL kernels' durations and where they sit relative to the deformation backward, idle gaps.
usage (GPU box): python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/timeline_dp.py [--overlap]"""
import argparse, os, sys, collections
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from torch.profiler import profile, ProfilerActivity

ap = argparse.ArgumentParser(); ap.add_argument("--overlap", action="store_true"); a = ap.parse_args()
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dev = f"cuda:{lr}"
os.dup2(2, 1)
dist.init_process_group("nccl", device_id=torch.device(dev))
import fdgs
from fdgs import _lib as L
from fdgs.render import render as render_fn
from fdgs.dp import train_iteration
from fdgs.losses import l1_view_loss as fused_loss
L.load()
cfg, args, pc, net, cams, gts_host = bench.build_scene("dynerf", "fdgs", dev, 0, 0)
bucket = bench.make_bucket(pc, net, dev)
bg = torch.ones(3, device=dev)
gts = [g.to(dev) for g in gts_host]
step = lambda: train_iteration(render_fn, pc, cams, gts, bg, bucket, rank, world, loss_fn=fused_loss, speculative=True, overlap_sh=a.overlap)
for _ in range(5):
    step()
torch.cuda.synchronize(); dist.barrier()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        step()
    torch.cuda.synchronize()
if rank == 0:
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    evs.sort(key=lambda e: e.time_range.start)
    t0, t1 = evs[0].time_range.start, max(e.time_range.end for e in evs)
    busy = collections.defaultdict(float); cnt = collections.Counter()
    for e in evs:
        key = e.name[:70]
        busy[key] += e.time_range.end - e.time_range.start; cnt[key] += 1
    # union of busy intervals (streams overlap)
    iv = sorted((e.time_range.start, e.time_range.end) for e in evs)
    cover, end = 0.0, iv[0][0]
    for s, f in iv:
        if f > end:
            cover += f - max(s, end); end = f
    print(f"world {world} overlap {a.overlap}: 3 iterations span {t1 - t0:.0f} us ({(t1 - t0) / 3:.0f} us / iteration), GPU busy (union) {cover:.0f} us, idle {100 * (1 - cover / (t1 - t0)):.1f} %")
    for k, v in sorted(busy.items(), key=lambda x: -x[1])[:30]:
        print(f"{v / 3:9.1f} us/iter  x{cnt[k] // 3:<3d} {k}")
    print("NCCL kernels (start relative to the iteration's first kernel, duration) and the deformation backward kernels:")
    for e in evs:
        if "nccl" in e.name.lower() or "k_deform_bwd" in e.name or "k_deform_sample" in e.name or "k_blend_bwd" in e.name:
            print(f"   {e.time_range.start - t0:9.0f} us  +{e.time_range.end - e.time_range.start:7.0f} us  {e.name[:60]}")
dist.barrier(); dist.destroy_process_group()
