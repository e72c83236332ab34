#!/bin/bash
# 8 GPUs: C2 weak scaling with the native driver (replicated and distributed array); then config C4
# (one index over 4 G chars, 500 M per GPU, properties verified on every rank)
set -u
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
nproc; free -g | head -2
( timeout 300 python -m pytest "tests/test_gpu_multi_native.py::test_class_devices_keyword" -x -q -m gpu ) > gpurun_out/r2_n8_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r2_n8_tests.log
timeout 600 $TR --master-port 29521 bench.py --gpus 8 --steps 5 --warmup 2 > gpurun_out/r2_bench_n8_native.json 2> gpurun_out/r2_bench_n8_native.err; echo "n8 rc=$?"
timeout 600 $TR --master-port 29523 bench.py --gpus 8 --steps 5 --warmup 2 --sharded-sa --no-cpu-baseline > gpurun_out/r2_bench_n8_sharded.json 2> gpurun_out/r2_bench_n8_sharded.err; echo "n8 sharded rc=$?"
timeout 1500 $TR --master-port 29522 bench.py --gpus 8 --steps 2 --warmup 1 --chars-per-gpu 500000000 --share-builder --verify-properties --no-cpu-baseline > gpurun_out/r2_bench_c4.json 2> gpurun_out/r2_bench_c4.err; echo "c4 rc=$?"
tail -3 gpurun_out/r2_bench_c4.err
