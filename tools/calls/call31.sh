#!/bin/bash
# final round-2 evidence on one B200: tests, smoke, both arms at config 4, fdgs arm at the other configs, ncu launch list + full captures, timeline
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/r02_gputest_final.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r02_gputest_final.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke_final.log 2>&1; echo "smoke rc=$?"
timeout 400 python bench.py --steps 10 --warmup 3 > gpurun_out/r02_bench_fdgs_dynerf.json 2> gpurun_out/r02_bench_fdgs_dynerf.err; echo "bench rc=$?"
timeout 400 python bench.py --impl reference --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_reference_dynerf.json 2> gpurun_out/r02_bench_reference_dynerf.err; echo "ref rc=$?"
for c in dnerf50k dnerf hypernerf; do
  timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --config $c > gpurun_out/r02_bench_fdgs_$c.json 2> gpurun_out/r02_bench_fdgs_$c.err; echo "$c rc=$?"
done
timeout 400 python bench.py --config infer1080p --mode render-sweep --steps 2 --warmup 3 > gpurun_out/r02_bench_fdgs_infer1080p_sweep.json 2> gpurun_out/r02_bench_fdgs_infer1080p_sweep.err; echo "sweep rc=$?"
timeout 200 python tools/timeline.py --views 8 > gpurun_out/r02_timeline_8views.log 2>&1; echo "timeline rc=$?"
CMD="python bench.py --views 1 --steps 1 --warmup 3 --no-extras --no-cpu-baseline"
timeout 200 $CMD > gpurun_out/r02_ncu_plain.log 2>&1 &&
timeout 500 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/r02_ncu_fdgs_launches_raw.csv $CMD > gpurun_out/r02_ncu_fdgs.log 2>&1
python tools/ncu_summary.py launches gpurun_out/r02_ncu_fdgs_launches_raw.csv > gpurun_out/r02_ncu_launches_summary.csv 2>&1
rm -f gpurun_out/r02_ncu_fdgs_launches_raw.csv
timeout 500 ncu --set full --clock-control none --import-source on -k regex:"k_deform_fwd_tc|k_deform_bwd|k_deform_sample|k_preprocess" --launch-skip 24 --launch-count 6 \
  -o gpurun_out/r02_full_final -f $CMD > gpurun_out/r02_ncu_full_final.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out/*.ncu-rep
