#!/bin/bash
set -u
mkdir -p gpurun_out
( time python -m pytest tests/test_gpu_query_semantics.py tests/test_gpu_dropin.py tests/test_gpu_dropin_cython.py tests/test_gpu_csv_columns.py tests/test_gpu_sharded.py -q -m gpu ) > gpurun_out/r2_pytest_gpu_b.log 2>&1; echo "pytest rc=$?"; tail -60 gpurun_out/r2_pytest_gpu_b.log
