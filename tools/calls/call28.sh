#!/bin/bash
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_gpu_raster.py tests/test_gpu_glue.py -m gpu -x -q > gpurun_out/r02_gputest_28.log 2>&1
rc=$?; echo "pytest rc=$rc" >> gpurun_out/r02_gputest_28.log
tail -3 gpurun_out/r02_gputest_28.log
if [ $rc -ne 0 ]; then exit 0; fi
B="timeout 200 python bench.py --steps 6 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/r02_ab_new.json 2> gpurun_out/r02_ab_new.err; echo "bench rc=$?"
FDGS_LIB=$PWD/tools/_bin/libfdgs_prev.so $B > gpurun_out/r02_ab_prev.json 2> gpurun_out/r02_ab_prev.err; echo "bench1 rc=$?"
