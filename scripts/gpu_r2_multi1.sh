#!/bin/bash
set -u
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests/test_gpu_multi_native.py -x -q -m gpu ) > gpurun_out/r2_multi_native.log 2>&1; echo "pytest rc=$?"; tail -40 gpurun_out/r2_multi_native.log
