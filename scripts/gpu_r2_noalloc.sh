#!/bin/bash
# the no-alloc entry point: alone in a fresh process, then at the end of the whole GPU suite; then smoke + bench
set -u
mkdir -p gpurun_out
( time python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu ) > gpurun_out/r2_noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -8 gpurun_out/r2_noalloc_alone.log
( time python -m pytest tests -x -q -m gpu ) > gpurun_out/r2_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/r2_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2_smoke.log
python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench_n1_base.json 2> gpurun_out/r2_bench_n1_base.err; echo "bench rc=$?"; cat gpurun_out/r2_bench_n1_base.json
