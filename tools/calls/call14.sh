#!/bin/bash
mkdir -p gpurun_out
# the allocation-free entry point alone in a fresh process, then the whole GPU suite (it runs last there), then smoke + a short bench
timeout 400 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q --timeout 300 > gpurun_out/r02_noalloc_alone.log 2>&1
echo "alone rc=$?" >> gpurun_out/r02_noalloc_alone.log
tail -3 gpurun_out/r02_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 600 > gpurun_out/r02_gputest_14.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_gputest_14.log
tail -4 gpurun_out/r02_gputest_14.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke_14.log 2>&1; echo "smoke rc=$?"
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_14.json 2> gpurun_out/r02_bench_14.err; echo "bench rc=$?"
timeout 240 compute-sanitizer --tool memcheck python -m pytest tests/test_zz_no_alloc.py -m gpu -q -k twice > gpurun_out/r02_noalloc_memcheck.log 2>&1
echo "memcheck rc=$?"; tail -4 gpurun_out/r02_noalloc_memcheck.log
