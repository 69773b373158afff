#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_deform.py tests/test_gpu_e2e.py tests/test_gpu_glue.py -m gpu -x -q --timeout 300 > gpurun_out/r02_gputest_15.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_gputest_15.log
tail -5 gpurun_out/r02_gputest_15.log
B="timeout 300 python bench.py --steps 8 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/r02_ab_outcc1.json 2> gpurun_out/r02_ab_outcc1.err; echo "bench rc=$?"
FDGS_LIB=$PWD/tools/_bin/libfdgs_outcc0.so $B > gpurun_out/r02_ab_outcc0.json 2> gpurun_out/r02_ab_outcc0.err; echo "bench0 rc=$?"
$B > gpurun_out/r02_ab_outcc1b.json 2> gpurun_out/r02_ab_outcc1b.err
