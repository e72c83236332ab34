#!/bin/bash
# sweeps with CPU arms (C1, C5, C3), the 2.3 GB CSV partition test, C4 CPU baselines, ncu launch list + --set full
set -u
mkdir -p gpurun_out
timeout 1500 python scripts/sweeps.py --out gpurun_out/r2_sweeps.json --sections c1,c5,c3 > gpurun_out/r2_sweeps.log 2>&1; echo "sweeps rc=$?"; tail -3 gpurun_out/r2_sweeps.log | cut -c1-600
SA_TEST_BIG_CSV=1 timeout 1200 python -m pytest tests/test_gpu_big_csv.py -x -q -m gpu > gpurun_out/r2_big_csv.log 2>&1; echo "big csv rc=$?"; tail -5 gpurun_out/r2_big_csv.log
timeout 1500 python scripts/cpu_c4.py --out gpurun_out/r2_cpu_c4.json > gpurun_out/r2_cpu_c4.log 2>&1; echo "cpu c4 rc=$?"; tail -1 gpurun_out/r2_cpu_c4.log | cut -c1-800
# ncu: launch list of the bench command (after a plain run), then --set full of the hot kernels
timeout 300 python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/r2_ncu_plain.log 2>&1; echo "plain rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/r2_ncu_list.log 2>&1; echo "ncu list rc=$?"
OS_SKIP=27 OTHER_SKIP=21 bash scripts/gpu_ncu_full.sh > gpurun_out/r2_ncu_full.log 2>&1; echo "ncu full rc=$?"; tail -4 gpurun_out/r2_ncu_full.log
