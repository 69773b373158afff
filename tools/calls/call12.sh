#!/bin/bash
mkdir -p gpurun_out
timeout 400 python bench.py --steps 10 --warmup 3 > gpurun_out/r02_bench_fdgs_dynerf_final.json 2> gpurun_out/r02_bench_fdgs_dynerf_final.err
timeout 400 python bench.py --impl reference --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_reference_dynerf_final.json 2> gpurun_out/r02_bench_reference_dynerf_final.err
timeout 200 python tools/timeline.py --views 8 > gpurun_out/r02_timeline_2.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/r02_ncu_fdgs_launches_raw.csv \
  python bench.py --views 1 --steps 1 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r02_ncu_fdgs.log 2>&1
python tools/ncu_summary.py launches gpurun_out/r02_ncu_fdgs_launches_raw.csv > gpurun_out/r02_ncu_fdgs_launches_summary.csv 2>&1
rm -f gpurun_out/r02_ncu_fdgs_launches_raw.csv
timeout 420 ncu --set full --clock-control none --import-source on -k regex:"k_tile_sort|k_scatter|k_blend|k_preprocess" --launch-skip 35 --launch-count 7 \
  -o gpurun_out/r02_full_raster -f python bench.py --views 1 --steps 1 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r02_ncu_full_raster.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_deform_fwd_tc|k_deform_bwd|k_deform_sample" --launch-skip 20 --launch-count 4 \
  -o gpurun_out/r02_full_deform -f python bench.py --views 1 --steps 1 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r02_ncu_full_deform.log 2>&1
ls -la gpurun_out/*.ncu-rep
