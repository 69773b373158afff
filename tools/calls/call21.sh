#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --views 1 --steps 1 --warmup 3 --no-extras --no-cpu-baseline"
timeout 200 $CMD > gpurun_out/r02_ncu_plain.log 2>&1 &&
timeout 400 ncu --set full --clock-control none --import-source on -k regex:"k_deform_bwd_weights|k_deform_bwd_data" --launch-skip 6 --launch-count 2 \
  -o gpurun_out/r02_full_bwd2 -f $CMD > gpurun_out/r02_ncu_full_bwd2.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/*.ncu-rep
