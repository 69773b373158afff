#!/bin/bash
mkdir -p gpurun_out
timeout 150 python -m pytest tests/test_gpu_deform.py tests/test_gpu_e2e.py tests/test_gpu_train_glue.py -m gpu -x -q > gpurun_out/r02_gputest_33.log 2>&1
rc=$?; echo "pytest rc=$rc" >> gpurun_out/r02_gputest_33.log; tail -2 gpurun_out/r02_gputest_33.log
if [ $rc -ne 0 ]; then exit 0; fi
for c in dnerf50k dnerf; do
  timeout 200 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extras --config $c > gpurun_out/r02_host_$c.json 2> gpurun_out/r02_host_$c.err; echo "$c rc=$?"
done
timeout 100 python tools/host_profile.py --views 1 > gpurun_out/r02_host_profile_1view.log 2>&1; head -3 gpurun_out/r02_host_profile_1view.log
