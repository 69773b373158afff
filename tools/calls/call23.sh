#!/bin/bash
mkdir -p gpurun_out
FDGS_LIB=$PWD/tools/_bin/libfdgs_ptbwd.so timeout 200 python bench.py --views 1 --steps 1 --warmup 1 --no-extras --no-cpu-baseline > gpurun_out/r02_pt_bwd2.log 2>&1
grep -c "bwd_weights cta" gpurun_out/r02_pt_bwd2.log
