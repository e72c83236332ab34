#!/bin/bash
# Wider onesweep digits, measured: builds scripts/bench_onesweep_wide here or on the box and runs it.
#   gpurun --timeout 600 -- 'bash scripts/gpu_onesweep_wide.sh'        (about 3 minutes of box time)
# Output: gpurun_out/onesweep_wide.log -- whole-sort times of the product pass (8 x 8 bits) against
# 7- and 6-pass digit plans, each run checked for order and stability ("ok" / "NOT SORTED").
set -u
mkdir -p gpurun_out
if [ ! -x scripts/bench_onesweep_wide ]; then
    nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -o scripts/bench_onesweep_wide \
        scripts/bench_onesweep_wide.cu || exit 1
fi
# a small run first: a wrong variant shows up in a second instead of hanging a 100 M-record sort
timeout 120 scripts/bench_onesweep_wide 3000000 1 > gpurun_out/onesweep_wide_small.log 2>&1; echo "small rc=$?"
grep -c " ok" gpurun_out/onesweep_wide_small.log; grep -v " ok" gpurun_out/onesweep_wide_small.log | head
timeout 400 scripts/bench_onesweep_wide 100000000 3 > gpurun_out/onesweep_wide.log 2>&1; echo "full rc=$?"
cat gpurun_out/onesweep_wide.log
