#!/bin/bash
# 2 GPUs: the native driver on real peers (in-process peer access; torchrun + shm + IPC), bench native vs python driver, reference arm under torchrun
set -u
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
( timeout 600 python -m pytest tests/test_gpu_multi_native.py "tests/test_gpu_dropin.py::test_two_devices_in_one_process" -x -q -m gpu ) > gpurun_out/r2_n2_tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/r2_n2_tests.log
timeout 600 $TR --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 2 --verify-single > gpurun_out/r2_bench_n2_native.json 2> gpurun_out/r2_bench_n2_native.err; echo "native rc=$?"
timeout 600 $TR --master-port 29512 bench.py --gpus 2 --steps 5 --warmup 2 --driver python --no-cpu-baseline > gpurun_out/r2_bench_n2_python.json 2> gpurun_out/r2_bench_n2_python.err; echo "python rc=$?"
timeout 600 $TR --master-port 29513 bench.py --gpus 2 --steps 2 --warmup 1 --impl reference > gpurun_out/r2_bench_n2_ref.json 2> gpurun_out/r2_bench_n2_ref.err; echo "ref rc=$?"
tail -3 gpurun_out/r2_bench_n2_native.err
