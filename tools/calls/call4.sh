#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/r02_parity_*.log
timeout 900 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/r02_gputest_4.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_gputest_4.log
tail -5 gpurun_out/r02_gputest_4.log
B="timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_fdgs_dynerf_2.json 2> gpurun_out/r02_bench_fdgs_dynerf_2.err
FDGS_ONE_SHOT=0 $B > gpurun_out/r02_ab_exact.json 2> gpurun_out/r02_ab_exact.err
for v in fwd_s1 bwd_s1 bwd_s1_old bwd_s2_m5 bwd_s2_m4b256; do
  FDGS_LIB=$PWD/tools/_bin/libfdgs_$v.so $B > gpurun_out/r02_ab_blend_$v.json 2> gpurun_out/r02_ab_blend_$v.err
done
FDGS_DW2_TENSOR=1 $B > gpurun_out/r02_ab_dw2_tensor.json 2> gpurun_out/r02_ab_dw2_tensor.err
FDGS_DW2_COST=0.25 $B > gpurun_out/r02_ab_dw2_c025.json 2> gpurun_out/r02_ab_dw2_c025.err
FDGS_DW2_COST=0.55 $B > gpurun_out/r02_ab_dw2_c055.json 2> gpurun_out/r02_ab_dw2_c055.err
for v in pt_fwd pt_bwd; do
  FDGS_LIB=$PWD/tools/_bin/libfdgs_$v.so timeout 200 python bench.py --views 1 --steps 1 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r02_$v.log 2>&1
done
timeout 200 python tools/timeline.py --views 8 > gpurun_out/r02_timeline_1.log 2>&1
timeout 200 python tools/host_profile.py --views 1 > gpurun_out/r02_host_profile_v1_1.log 2>&1
timeout 420 ncu --set full --clock-control none --import-source on -k regex:"k_tile_sort|k_scatter|k_blend|k_preprocess" --launch-skip 35 --launch-count 7 \
  -o gpurun_out/r02_full_raster python bench.py --views 1 --steps 1 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r02_ncu_full_raster.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_deform_fwd_tc|k_deform_bwd|k_deform_sample" --launch-skip 20 --launch-count 4 \
  -o gpurun_out/r02_full_deform python bench.py --views 1 --steps 1 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r02_ncu_full_deform.log 2>&1
ls -la gpurun_out/*.ncu-rep
