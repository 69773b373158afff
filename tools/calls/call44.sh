#!/bin/bash
mkdir -p gpurun_out
timeout 90 python bench.py --steps 10 --warmup 3 > gpurun_out/r02_bench_fdgs_dynerf.json 2> gpurun_out/r02_bench_fdgs_dynerf.err; echo "bench rc=$?"
timeout 60 python bench.py --impl reference --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_reference_dynerf.json 2> gpurun_out/r02_bench_reference_dynerf.err; echo "ref rc=$?"
