#!/bin/bash
mkdir -p gpurun_out
timeout 240 python -m pytest tests/test_gpu_deform.py tests/test_gpu_e2e.py -m gpu -x -q > gpurun_out/r02_gputest_39.log 2>&1
rc=$?; echo "pytest rc=$rc" >> gpurun_out/r02_gputest_39.log; tail -3 gpurun_out/r02_gputest_39.log
if [ $rc -ne 0 ]; then exit 0; fi
for fl in 32 0 64; do
  FDGS_W_FLUSH=$fl FDGS_TEST_P=300000 timeout 200 python -m pytest tests/test_gpu_deform.py -m gpu -q -k "baseline_shapes and dynerf" > gpurun_out/r02_gputest_39_$fl.log 2>&1
  echo "300k flush=$fl rc=$?"; grep -h "AssertionError:\|passed\|failed" gpurun_out/r02_gputest_39_$fl.log | tail -2; grep "baseline_shapes\[dynerf\]: worst" gpurun_out/r02_parity_deform.log | tail -1
done
B="timeout 200 python bench.py --steps 6 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/r02_ab_fl32.json 2> gpurun_out/r02_ab_fl32.err; echo "rc=$?"
FDGS_W_FLUSH=0 $B > gpurun_out/r02_ab_fl0.json 2> gpurun_out/r02_ab_fl0.err; echo "rc=$?"
FDGS_W_FLUSH=64 $B > gpurun_out/r02_ab_fl64.json 2> gpurun_out/r02_ab_fl64.err; echo "rc=$?"
