#!/bin/bash
mkdir -p gpurun_out
export FDGS_LIB=$PWD/tools/_bin/libfdgs_blockmap.so
timeout 25 python -m pytest tests/test_gpu_deform.py -m gpu -x -q > gpurun_out/r02_gputest_46.log 2>&1; echo "pytest rc=$?"; tail -1 gpurun_out/r02_gputest_46.log
timeout 20 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/r02_ab_blockmap.json 2> gpurun_out/r02_ab_blockmap.err; echo "bench rc=$?"
