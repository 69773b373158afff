#!/bin/bash
mkdir -p gpurun_out
N=8
R="timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513"
$R tools/timeline_dp.py > gpurun_out/r02_timeline_dp8.log 2>&1
$R tools/timeline_dp.py --overlap > gpurun_out/r02_timeline_dp8_overlap.log 2>&1
T="$R bench.py --gpus $N --steps 30 --warmup 5 --no-cpu-baseline --no-extras"
FDGS_OVERLAP_SH=1 FDGS_OVERLAP_SPARE_SMS=20 $T > gpurun_out/r02_bench_fdgs_8gpu_overlap_s20.json 2> gpurun_out/r02_bench_fdgs_8gpu_overlap_s20.err
FDGS_OVERLAP_SH=1 FDGS_OVERLAP_SPARE_SMS=0 $T > gpurun_out/r02_bench_fdgs_8gpu_overlap_s0.json 2> gpurun_out/r02_bench_fdgs_8gpu_overlap_s0.err
FDGS_OVERLAP_SH=1 NCCL_MAX_NCHANNELS=8 $T > gpurun_out/r02_bench_fdgs_8gpu_overlap_c8.json 2> gpurun_out/r02_bench_fdgs_8gpu_overlap_c8.err
grep -h "iterations span" gpurun_out/r02_timeline_dp8*.log
