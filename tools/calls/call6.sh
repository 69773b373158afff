#!/bin/bash
# 2-GPU: scaling with / without the early all-reduce, dW2 cost A/B on rank 0's kernels
mkdir -p gpurun_out
T="timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu-baseline"
$T > gpurun_out/r02_bench_fdgs_2gpu.json 2> gpurun_out/r02_bench_fdgs_2gpu.err
FDGS_OVERLAP_SH=1 $T > gpurun_out/r02_bench_fdgs_2gpu_overlap.json 2> gpurun_out/r02_bench_fdgs_2gpu_overlap.err
FDGS_ONE_SHOT=0 $T > gpurun_out/r02_bench_fdgs_2gpu_exact.json 2> gpurun_out/r02_bench_fdgs_2gpu_exact.err
for c in 0.55 0.65; do
  FDGS_DW2_COST=$c timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02_ab_dw2_c$c.json 2> gpurun_out/r02_ab_dw2_c$c.err
done
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_fdgs_dynerf_4.json 2> gpurun_out/r02_bench_fdgs_dynerf_4.err
tail -2 gpurun_out/r02_bench_fdgs_2gpu.err
