#!/bin/bash
mkdir -p gpurun_out
B="timeout 200 python bench.py --steps 6 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/r02_ab_c0.json 2> gpurun_out/r02_ab_c0.err; echo "rc=$?"
FDGS_W_COSTS=0.63,0.96,0.94 $B > gpurun_out/r02_ab_c1.json 2> gpurun_out/r02_ab_c1.err; echo "rc=$?"
FDGS_W_COSTS=0.60,1.0,1.0 $B > gpurun_out/r02_ab_c2.json 2> gpurun_out/r02_ab_c2.err; echo "rc=$?"
FDGS_W_COSTS=0.70,0.96,0.94 $B > gpurun_out/r02_ab_c3.json 2> gpurun_out/r02_ab_c3.err; echo "rc=$?"
