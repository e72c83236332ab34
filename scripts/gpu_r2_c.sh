#!/bin/bash
set -u
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -x -q -m gpu ) > gpurun_out/r2_pytest_gpu_c.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r2_pytest_gpu_c.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench_n1_c.json 2> gpurun_out/r2_bench_n1_c.err; echo "bench rc=$?"
tail -2 gpurun_out/r2_bench_n1_c.err
