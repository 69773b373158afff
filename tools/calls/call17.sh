#!/bin/bash
mkdir -p gpurun_out
timeout 150 python -m pytest tests/test_gpu_deform.py tests/test_gpu_e2e.py tests/test_gpu_glue.py -m gpu -x -q --timeout 120 > gpurun_out/r02_gputest_17.log 2>&1
rc=$?; echo "pytest rc=$rc" >> gpurun_out/r02_gputest_17.log
tail -5 gpurun_out/r02_gputest_17.log
if [ $rc -ne 0 ]; then exit 0; fi
B="timeout 300 python bench.py --steps 8 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/r02_ab_smp1.json 2> gpurun_out/r02_ab_smp1.err; echo "bench rc=$?"
FDGS_LIB=$PWD/tools/_bin/libfdgs_nosmp.so $B > gpurun_out/r02_ab_smp0.json 2> gpurun_out/r02_ab_smp0.err; echo "bench0 rc=$?"
FDGS_LIB=$PWD/tools/_bin/libfdgs_ptfwd.so timeout 300 python bench.py --views 1 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r02_pt_fwd3.log 2>&1
grep "fwd phases" gpurun_out/r02_pt_fwd3.log | tail -2
