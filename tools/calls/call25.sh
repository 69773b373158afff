#!/bin/bash
mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_gpu_deform.py tests/test_gpu_e2e.py tests/test_gpu_glue.py -m gpu -x -q > gpurun_out/r02_gputest_25.log 2>&1
rc=$?; echo "pytest rc=$rc" >> gpurun_out/r02_gputest_25.log
tail -5 gpurun_out/r02_gputest_25.log
if [ $rc -ne 0 ]; then exit 0; fi
B="timeout 200 python bench.py --steps 8 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/r02_ab_bepi16.json 2> gpurun_out/r02_ab_bepi16.err; echo "bench rc=$?"
FDGS_LIB=$PWD/tools/_bin/libfdgs_bepi8.so $B > gpurun_out/r02_ab_bepi8.json 2> gpurun_out/r02_ab_bepi8.err; echo "bench0 rc=$?"
