#!/bin/bash
mkdir -p gpurun_out
B="timeout 50 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-extras"
FDGS_TEST_P=300000 timeout 40 python -m pytest tests/test_gpu_deform.py -m gpu -q -x -k "baseline_shapes and dynerf" > gpurun_out/r02_gputest_43a.log 2>&1
rc=$?; echo "300k v1 rc=$rc"; grep -h "AssertionError:\|passed\|failed" gpurun_out/r02_gputest_43a.log | tail -1
if [ $rc -ne 0 ]; then exit 0; fi
$B > gpurun_out/r02_ab_d1.json 2> gpurun_out/r02_ab_d1.err; echo "rc=$?"
FDGS_W_FLUSH=0 $B > gpurun_out/r02_ab_d1_fl0.json 2> gpurun_out/r02_ab_d1_fl0.err; echo "rc=$?"
FDGS_LIB=$PWD/tools/_bin/libfdgs_drain3.so $B > gpurun_out/r02_ab_d3.json 2> gpurun_out/r02_ab_d3.err; echo "rc=$?"
timeout 60 python -m pytest tests/test_gpu_deform.py tests/test_gpu_e2e.py -m gpu -x -q > gpurun_out/r02_gputest_43.log 2>&1; echo "pytest rc=$?"; tail -1 gpurun_out/r02_gputest_43.log
