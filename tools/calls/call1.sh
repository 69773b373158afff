#!/bin/bash
# GPU call 1 (round 2): parity suite, both bench arms, launch lists of both arms, timeline
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x --timeout 900 > gpurun_out/r02_gputest_1.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_gputest_1.log
python bench.py --impl reference --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_reference_dynerf_0.json 2> gpurun_out/r02_bench_reference_dynerf_0.err
python bench.py --steps 10 --warmup 3 > gpurun_out/r02_bench_fdgs_dynerf_0.json 2> gpurun_out/r02_bench_fdgs_dynerf_0.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file gpurun_out/r02_ncu_reference_launches_raw.csv \
  python bench.py --impl reference --views 1 --steps 1 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r02_ncu_reference.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/r02_ncu_fdgs_launches_raw.csv \
  python bench.py --views 1 --steps 1 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r02_ncu_fdgs.log 2>&1
python tools/timeline.py --views 8 > gpurun_out/r02_timeline_0.log 2>&1
python tools/ncu_summary.py launches gpurun_out/r02_ncu_reference_launches_raw.csv > gpurun_out/r02_ncu_reference_launches_summary.csv 2>&1
python tools/ncu_summary.py launches gpurun_out/r02_ncu_fdgs_launches_raw.csv > gpurun_out/r02_ncu_fdgs_launches_summary.csv 2>&1
rm -f gpurun_out/r02_ncu_reference_launches_raw.csv gpurun_out/r02_ncu_fdgs_launches_raw.csv
tail -5 gpurun_out/r02_gputest_1.log
