#!/bin/bash
# One gpurun call: GPU parity tests (all of them, no -x: every failure is wanted), the full-size
# check of an index with more than 2^31 entries, then the default bench line.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
nproc > gpurun_out/host.txt; free -g >> gpurun_out/host.txt
timeout 1200 python -m pytest tests -m gpu -q --durations=15 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 600 python scripts/check_wide.py --out gpurun_out/wide.json > gpurun_out/wide.log 2>&1; echo "wide rc=$?" | tee -a gpurun_out/wide.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
tail -25 gpurun_out/pytest_gpu.log; tail -5 gpurun_out/wide.log; cat gpurun_out/bench.json | cut -c1-600; tail -3 gpurun_out/bench.err
