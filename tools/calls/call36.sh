#!/bin/bash
mkdir -p gpurun_out
N=2
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/r02_bench_fdgs_2gpu.json 2> gpurun_out/r02_bench_fdgs_2gpu.err
echo "rc=$?"
