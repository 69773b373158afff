#!/bin/bash
mkdir -p gpurun_out
for fl in 0 32; do
FDGS_W_FLUSH=$fl FDGS_LIB=$PWD/tools/_bin/libfdgs_ptbwd.so timeout 200 python bench.py --views 1 --steps 1 --warmup 1 --no-extras --no-cpu-baseline > gpurun_out/r02_pt_bwd_fl$fl.log 2>&1
done
