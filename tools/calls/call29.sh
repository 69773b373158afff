#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --views 1 --steps 1 --warmup 3 --no-extras --no-cpu-baseline"
timeout 200 $CMD > gpurun_out/r02_ncu_plain.log 2>&1 &&
timeout 400 ncu --set full --clock-control none --import-source on -k regex:"k_blend_bwd" --launch-skip 3 --launch-count 1 \
  -o gpurun_out/r02_full_blendbwd -f $CMD > gpurun_out/r02_ncu_full_blendbwd.log 2>&1
echo "ncu rc=$?"
