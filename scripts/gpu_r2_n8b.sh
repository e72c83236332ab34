#!/bin/bash
# 8 GPUs after the asynchronous text all-gather: replicated array (verified against the single-GPU build), then the distributed array
set -u
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
SA_BENCH_DEBUG="timeline=1" timeout 600 $TR --master-port 29531 bench.py --gpus 8 --steps 5 --warmup 3 --no-cpu-baseline --verify-single > gpurun_out/r2_bench_n8_async.json 2> gpurun_out/r2_bench_n8_async.err; echo "n8 rc=$?"
grep "timeline rank 0" gpurun_out/r2_bench_n8_async.err | tail -1
python - <<'PY'
import json
d = json.load(open("gpurun_out/r2_bench_n8_async.json"))
print("replicated build", d["build_ms_per_step"], "e2e", d["e2e"]["build_ms"], "qk", d.get("query_kernel_ms"), d["queries_per_s"], d["record_queries_per_s"], d.get("verify"))
print(d["multi_gpu"]["phases_ms_max_over_ranks"]); print(d["multi_gpu"]["phases_ms_by_rank"])
PY
timeout 600 $TR --master-port 29532 bench.py --gpus 8 --steps 5 --warmup 3 --no-cpu-baseline --sharded-sa --verify-single > gpurun_out/r2_bench_n8_sharded.json 2> gpurun_out/r2_bench_n8_sharded.err; echo "n8 sharded rc=$?"
python - <<'PY'
import json
d = json.load(open("gpurun_out/r2_bench_n8_sharded.json"))
print("sharded build", d["build_ms_per_step"], "e2e", d["e2e"]["build_ms"], "qk", d.get("query_kernel_ms"), d["queries_per_s"], d["record_queries_per_s"], d.get("verify"))
print(d["multi_gpu"]["phases_ms_max_over_ranks"])
PY
