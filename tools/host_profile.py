"""Host-side cost of one training iteration: wall time until every launch is queued (no sync) vs GPU time, and a cProfile
of the Python glue.  usage (GPU box): python tools/host_profile.py [--views 8]"""
import argparse, cProfile, os, pstats, sys, time, io
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
ap = argparse.ArgumentParser(); ap.add_argument("--views", type=int, default=8); a = ap.parse_args()
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
step = lambda: train_iteration(render_fn, pc, cams, gts, bg, bucket, 0, 1, loss_fn=fused_loss, speculative=True)
for _ in range(3):
    step()
torch.cuda.synchronize()
for rep in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter(); e0.record()
    step()
    e1.record(); t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print(f"views {a.views}: host queued everything after {1e3 * (t1 - t0):.2f} ms; GPU span {e0.elapsed_time(e1):.2f} ms; wall to idle {1e3 * (t2 - t0):.2f} ms")
pr = cProfile.Profile()
torch.cuda.synchronize()
N = 100                                  # pstats prints milliseconds: profile N iterations so that per-call costs show
pr.enable()
for _ in range(N):
    step()
    torch.cuda.synchronize()             # (outside the per-function times that matter: it is its own entry)
pr.disable()
print(f"cProfile over {N} iterations of {a.views} view(s): divide the times by {N}")
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(50); print(s.getvalue()[:11000])
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(40); print(s.getvalue()[:9000])
