#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_raster.py tests/test_gpu_e2e.py -m gpu -q --timeout 300 > gpurun_out/r02_gputest_13.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_gputest_13.log
tail -3 gpurun_out/r02_gputest_13.log
B="timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/r02_ab_sp3.json 2> gpurun_out/r02_ab_sp3.err
for v in sp0 sp1 sp2 sp6; do FDGS_LIB=$PWD/tools/_bin/libfdgs_$v.so $B > gpurun_out/r02_ab_$v.json 2> gpurun_out/r02_ab_$v.err; done
