#!/bin/bash
mkdir -p gpurun_out
FDGS_TEST_P=300000 timeout 200 python -m pytest tests/test_gpu_deform.py -m gpu -q -x -k "baseline_shapes and dynerf" > gpurun_out/r02_gputest_41.log 2>&1
echo "300k rc=$?"; grep -h "AssertionError:\|passed\|failed" gpurun_out/r02_gputest_41.log | tail -2
for fl in 0 32; do
FDGS_W_FLUSH=$fl FDGS_LIB=$PWD/tools/_bin/libfdgs_ptbwd.so timeout 200 python bench.py --views 1 --steps 1 --warmup 1 --no-extras --no-cpu-baseline > gpurun_out/r02_pt_bwd_fl$fl.log 2>&1
done
B="timeout 200 python bench.py --steps 6 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/r02_ab_fl32.json 2> gpurun_out/r02_ab_fl32.err; echo "rc=$?"
FDGS_W_FLUSH=0 $B > gpurun_out/r02_ab_fl0.json 2> gpurun_out/r02_ab_fl0.err; echo "rc=$?"
