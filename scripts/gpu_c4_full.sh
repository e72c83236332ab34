#!/bin/bash
# BASELINE config 4 at full size: ONE index over 4 G characters on 8 GPUs (500 M characters per GPU),
# verified by the size-independent properties (permutation, first-8-byte key order, replica digests),
# plus the default-size N=8 line.  Run as:  gpurun --gpus 8 --timeout 900 -- 'bash scripts/gpu_c4_full.sh'
# (charged 8x; the 2.2 G / 2-GPU run of the same command took 45 s of box time).
set -u
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531"
timeout 600 $TR bench.py --gpus 8 --steps 3 --warmup 3 --chars-per-gpu 500000000 --share-builder --verify-properties \
    > gpurun_out/bench_n8_c4.json 2> gpurun_out/bench_n8_c4.err; echo "c4 rc=$?"
cut -c1-600 gpurun_out/bench_n8_c4.json
timeout 300 $TR bench.py --gpus 8 > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err; echo "n8 rc=$?"
cut -c1-400 gpurun_out/bench_n8.json
