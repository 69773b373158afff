#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/r02_parity_*.log
python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/r02_gputest_2.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_gputest_2.log
python bench.py --impl reference --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_reference_dynerf_0.json 2> gpurun_out/r02_bench_reference_dynerf_0.err
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_fdgs_dynerf_1.json 2> gpurun_out/r02_bench_fdgs_dynerf_1.err
for v in b128 b128m8 b64m8 b128m6; do
  FDGS_LIB=$PWD/tools/_bin/libfdgs_$v.so python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02_ab_blendbwd_$v.json 2> gpurun_out/r02_ab_blendbwd_$v.err
done
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file gpurun_out/r02_ncu_reference_launches_raw.csv \
  python bench.py --impl reference --views 1 --steps 1 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r02_ncu_reference.log 2>&1
python tools/ncu_summary.py launches gpurun_out/r02_ncu_reference_launches_raw.csv > gpurun_out/r02_ncu_reference_launches_summary.csv 2>&1
rm -f gpurun_out/r02_ncu_reference_launches_raw.csv
python tools/host_profile.py --views 8 > gpurun_out/r02_host_profile_0.log 2>&1
python tools/host_profile.py --views 1 > gpurun_out/r02_host_profile_v1_0.log 2>&1
tail -8 gpurun_out/r02_gputest_2.log
