#!/bin/bash
mkdir -p gpurun_out
timeout 240 python -m pytest tests/test_gpu_deform.py -m gpu -x -q -k many_tiles > gpurun_out/r02_gputest_38.log 2>&1
rc=$?; echo "pytest rc=$rc" >> gpurun_out/r02_gputest_38.log; tail -3 gpurun_out/r02_gputest_38.log
FDGS_TEST_P=300000 timeout 200 python -m pytest tests/test_gpu_deform.py -m gpu -q -k "baseline_shapes and dynerf" > gpurun_out/r02_gputest_38b.log 2>&1
echo "300k rc=$?"; tail -3 gpurun_out/r02_gputest_38b.log; grep "baseline_shapes\[dynerf\]" gpurun_out/r02_parity_deform.log | tail -2
