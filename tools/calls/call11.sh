#!/bin/bash
mkdir -p gpurun_out
N=8
T="timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus $N --steps 30 --warmup 5 --no-cpu-baseline --no-extras"
$T > gpurun_out/r02_nccl_default.json 2> gpurun_out/r02_nccl_default.err
NCCL_PROTO=Simple $T > gpurun_out/r02_nccl_simple.json 2> gpurun_out/r02_nccl_simple.err
NCCL_ALGO=NVLS $T > gpurun_out/r02_nccl_nvls.json 2> gpurun_out/r02_nccl_nvls.err
NCCL_ALGO=Tree NCCL_PROTO=Simple $T > gpurun_out/r02_nccl_tree.json 2> gpurun_out/r02_nccl_tree.err
NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT,TUNING $T 2>&1 | grep -iE "nvls|algo|proto|channels" | head -30 > gpurun_out/r02_nccl_info.log
