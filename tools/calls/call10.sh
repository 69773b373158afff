#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/r02_parity_*.log
timeout 900 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/r02_gputest_10.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_gputest_10.log
tail -4 gpurun_out/r02_gputest_10.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke.log 2>&1; tail -1 gpurun_out/r02_smoke.log
for c in dnerf dynerf; do
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --config $c > gpurun_out/r02_ab2_epi16_$c.json 2> gpurun_out/r02_ab2_epi16_$c.err
FDGS_LIB=$PWD/tools/_bin/libfdgs_epi8.so timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --config $c > gpurun_out/r02_ab2_epi8_$c.json 2> gpurun_out/r02_ab2_epi8_$c.err
done
