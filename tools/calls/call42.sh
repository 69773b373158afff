#!/bin/bash
mkdir -p gpurun_out
FDGS_TEST_P=300000 timeout 45 python -m pytest tests/test_gpu_deform.py -m gpu -q -x -k "baseline_shapes and dynerf" > gpurun_out/r02_gputest_42a.log 2>&1
rc=$?; echo "300k rc=$rc"; grep -h "AssertionError:\|passed\|failed" gpurun_out/r02_gputest_42a.log | tail -2
if [ $rc -ne 0 ]; then exit 0; fi
timeout 90 python -m pytest tests/test_gpu_deform.py tests/test_gpu_e2e.py -m gpu -x -q > gpurun_out/r02_gputest_42.log 2>&1
rc=$?; echo "pytest rc=$rc"; tail -2 gpurun_out/r02_gputest_42.log
if [ $rc -ne 0 ]; then exit 0; fi
B="timeout 60 python bench.py --steps 6 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/r02_ab_fl32.json 2> gpurun_out/r02_ab_fl32.err; echo "rc=$?"
FDGS_W_FLUSH=0 $B > gpurun_out/r02_ab_fl0.json 2> gpurun_out/r02_ab_fl0.err; echo "rc=$?"
