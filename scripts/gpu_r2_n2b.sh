#!/bin/bash
# 2 GPUs: replicated vs distributed array (SA_FLAG_SHARDED_SA), the devices= class test on real peers
set -u
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
( timeout 600 python -m pytest tests/test_gpu_multi_native.py -x -q -m gpu ) > gpurun_out/r2_n2_tests_b.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/r2_n2_tests_b.log
timeout 600 $TR --master-port 29531 bench.py --gpus 2 --steps 5 --warmup 2 --verify-single --no-cpu-baseline > gpurun_out/r2_bench_n2_repl.json 2> gpurun_out/r2_bench_n2_repl.err; echo "repl rc=$?"
timeout 600 $TR --master-port 29532 bench.py --gpus 2 --steps 5 --warmup 2 --verify-single --no-cpu-baseline --sharded-sa > gpurun_out/r2_bench_n2_sharded.json 2> gpurun_out/r2_bench_n2_sharded.err; echo "sharded rc=$?"
tail -3 gpurun_out/r2_bench_n2_sharded.err
