#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/r02_parity_*.log
timeout 900 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/r02_gputest_5.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_gputest_5.log
tail -4 gpurun_out/r02_gputest_5.log
B="timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
timeout 400 python bench.py --steps 10 --warmup 3 > gpurun_out/r02_bench_fdgs_dynerf_3.json 2> gpurun_out/r02_bench_fdgs_dynerf_3.err
for v in relu_u2 relu_u4; do
  FDGS_LIB=$PWD/tools/_bin/libfdgs_$v.so $B > gpurun_out/r02_ab_$v.json 2> gpurun_out/r02_ab_$v.err
done
FDGS_LIB=$PWD/tools/_bin/libfdgs_pt_fwd.so timeout 200 python bench.py --views 1 --steps 1 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r02_pt_fwd.log 2>&1
timeout 200 python tools/timeline.py --views 8 > gpurun_out/r02_timeline_1.log 2>&1
timeout 200 python tools/host_profile.py --views 1 > gpurun_out/r02_host_profile_v1_1.log 2>&1
# every BASELINE config, both arms
for c in dnerf50k dnerf hypernerf; do
  timeout 300 python bench.py --config $c --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_fdgs_$c.json 2> gpurun_out/r02_bench_fdgs_$c.err
  timeout 400 python bench.py --config $c --impl reference --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_reference_$c.json 2> gpurun_out/r02_bench_reference_$c.err
done
timeout 400 python bench.py --impl reference --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_reference_dynerf.json 2> gpurun_out/r02_bench_reference_dynerf.err
timeout 400 python bench.py --config infer1080p --mode render-sweep --steps 2 --warmup 3 > gpurun_out/r02_bench_fdgs_infer1080p_sweep.json 2> gpurun_out/r02_bench_fdgs_infer1080p_sweep.err
timeout 600 python bench.py --config infer1080p --mode render-sweep --impl reference --steps 1 --warmup 3 > gpurun_out/r02_bench_reference_infer1080p_sweep.json 2> gpurun_out/r02_bench_reference_infer1080p_sweep.err
ls -la gpurun_out/*.json | tail -20
