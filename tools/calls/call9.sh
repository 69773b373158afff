#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_deform.py tests/test_gpu_e2e.py -m gpu -q --timeout 300 > gpurun_out/r02_gputest_9.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_gputest_9.log
tail -4 gpurun_out/r02_gputest_9.log
B="timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/r02_ab_epi16.json 2> gpurun_out/r02_ab_epi16.err
FDGS_LIB=$PWD/tools/_bin/libfdgs_epi8.so $B > gpurun_out/r02_ab_epi8.json 2> gpurun_out/r02_ab_epi8.err
$B --config hypernerf > gpurun_out/r02_ab_epi16_hypernerf.json 2> gpurun_out/r02_ab_epi16_hypernerf.err
FDGS_LIB=$PWD/tools/_bin/libfdgs_epi8.so $B --config hypernerf > gpurun_out/r02_ab_epi8_hypernerf.json 2> gpurun_out/r02_ab_epi8_hypernerf.err
FDGS_LIB=$PWD/tools/_bin/libfdgs_pt_fwd.so timeout 200 python bench.py --views 1 --steps 1 --warmup 3 --no-extras --no-cpu-baseline 2>&1 | grep "fwd phases" | tail -2 > gpurun_out/r02_pt_fwd_epi16.log
