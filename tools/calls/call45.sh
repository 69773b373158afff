#!/bin/bash
mkdir -p gpurun_out
timeout 70 python -m pytest tests -m gpu -x -q > gpurun_out/r02_gputest_final.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r02_gputest_final.log; tail -3 gpurun_out/r02_gputest_final.log
