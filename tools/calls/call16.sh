#!/bin/bash
mkdir -p gpurun_out
FDGS_LIB=$PWD/tools/_bin/libfdgs_ptfwd.so timeout 300 python bench.py --views 1 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r02_pt_fwd2.log 2>&1
grep "fwd phases" gpurun_out/r02_pt_fwd2.log | tail -3
