#!/bin/bash
# 8 GPUs: C2 weak scaling with the native driver (replicated array), then config C4
# (one index over 4 G chars, 500 M per GPU, properties verified on every rank)
set -u
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
nproc; free -g | head -2
timeout 600 $TR --master-port 29521 bench.py --gpus 8 --steps 8 --warmup 3 > gpurun_out/r2_bench_n8_native.json 2> gpurun_out/r2_bench_n8_native.err; echo "n8 rc=$?"
tail -c 1500 gpurun_out/r2_bench_n8_native.json
timeout 1500 $TR --master-port 29522 bench.py --gpus 8 --steps 2 --warmup 1 --chars-per-gpu 500000000 --share-builder --verify-properties --no-cpu-baseline > gpurun_out/r2_bench_c4.json 2> gpurun_out/r2_bench_c4.err; echo "c4 rc=$?"
tail -3 gpurun_out/r2_bench_c4.err
tail -c 1500 gpurun_out/r2_bench_c4.json
