#!/bin/bash
mkdir -p gpurun_out
timeout 240 python -m pytest tests/test_gpu_deform.py -m gpu -x -q -k many_tiles > gpurun_out/r02_gputest_37.log 2>&1
rc=$?; echo "pytest rc=$rc" >> gpurun_out/r02_gputest_37.log; tail -4 gpurun_out/r02_gputest_37.log
