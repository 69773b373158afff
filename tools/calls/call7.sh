#!/bin/bash
mkdir -p gpurun_out
N=$1
T="timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 30 --warmup 5 --no-cpu-baseline"
$T > gpurun_out/r02_bench_fdgs_${N}gpu.json 2> gpurun_out/r02_bench_fdgs_${N}gpu.err
FDGS_OVERLAP_SH=1 $T > gpurun_out/r02_bench_fdgs_${N}gpu_overlap.json 2> gpurun_out/r02_bench_fdgs_${N}gpu_overlap.err
FDGS_ONE_SHOT=0 $T > gpurun_out/r02_bench_fdgs_${N}gpu_exact.json 2> gpurun_out/r02_bench_fdgs_${N}gpu_exact.err
tail -2 gpurun_out/r02_bench_fdgs_${N}gpu.err
