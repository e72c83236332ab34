#!/bin/bash
# round 2, first GPU call: config-size parity tests, whole GPU suite, wide-digit experiment, bench both arms
set -u
mkdir -p gpurun_out
( time python -m pytest tests/test_gpu_config_parity.py -x -q -m gpu ) > gpurun_out/r2_config_parity.log 2>&1; echo "config parity rc=$?"; tail -5 gpurun_out/r2_config_parity.log
( time python -m pytest tests -x -q -m gpu --deselect tests/test_gpu_config_parity.py ) > gpurun_out/r2_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_pytest_gpu.log
bash scripts/gpu_onesweep_wide.sh > gpurun_out/r2_wide_driver.log 2>&1; echo "wide rc=$?"
python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench_n1_a.json 2> gpurun_out/r2_bench_n1_a.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_ref_a.json 2> gpurun_out/r2_bench_ref_a.err; echo "ref rc=$?"
nproc; free -g | head -2
