#!/bin/bash
mkdir -p gpurun_out
timeout 240 python -m pytest tests -m gpu -x -q > gpurun_out/r02_gputest_35.log 2>&1
rc=$?; echo "pytest rc=$rc" >> gpurun_out/r02_gputest_35.log; tail -2 gpurun_out/r02_gputest_35.log
if [ $rc -ne 0 ]; then exit 0; fi
for c in dnerf50k dnerf; do
  timeout 200 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extras --config $c > gpurun_out/r02_host_$c.json 2> gpurun_out/r02_host_$c.err; echo "$c rc=$?"
done
timeout 200 python tools/host_profile.py --views 1 > gpurun_out/r02_host_profile_1view.log 2>&1; head -3 gpurun_out/r02_host_profile_1view.log
