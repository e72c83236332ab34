#!/bin/bash
# sharded-SA mode on one GPU (emulated ranks), then the C3 record comparison again
set -u
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_multi_native.py tests/test_gpu_query_parity.py -x -q -m gpu ) > gpurun_out/r2_sharded_sa_tests.log 2>&1; echo "tests rc=$?"; tail -8 gpurun_out/r2_sharded_sa_tests.log
timeout 1500 python scripts/sweeps.py --out gpurun_out/r2_sweeps_c3.json --sections c3 > gpurun_out/r2_sweeps_c3.log 2>&1; echo "sweeps c3 rc=$?"; tail -2 gpurun_out/r2_sweeps_c3.log | cut -c1-200
