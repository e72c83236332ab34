#!/bin/bash
# C3 sweep (CSV end to end with the reference's CPU arm), the 2.3 GB CSV partition test, new load test
set -u
mkdir -p gpurun_out
timeout 1500 python scripts/sweeps.py --out gpurun_out/r2_sweeps_c3.json --sections c3 > gpurun_out/r2_sweeps_c3.log 2>&1; echo "sweeps c3 rc=$?"; tail -2 gpurun_out/r2_sweeps_c3.log | cut -c1-300
SA_TEST_BIG_CSV=1 timeout 1500 python -m pytest tests/test_gpu_big_csv.py tests/test_gpu_dropin.py -x -q -m gpu > gpurun_out/r2_big_csv.log 2>&1; echo "big csv rc=$?"; tail -5 gpurun_out/r2_big_csv.log
gcc -Iinclude tests/c_abi_smoke.c -Lsuffixarray_b200 -lsa_cuda -Wl,-rpath,$PWD/suffixarray_b200 -o /tmp/c_abi_smoke && /tmp/c_abi_smoke
