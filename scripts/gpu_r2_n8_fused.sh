#!/bin/bash
# 8 GPUs, one short line with the fused finish + replicate (the last GPU-minutes of the round)
mkdir -p gpurun_out
timeout 50 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus 8 --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench_n8_fused.json 2> gpurun_out/r2_bench_n8_fused.err; echo "rc=$?"
python - <<'PY'
import json
d = json.load(open("gpurun_out/r2_bench_n8_fused.json"))
print("build", d["build_ms_per_step"], "e2e", d["e2e"]["build_ms"], d["multi_gpu"]["phases_ms_max_over_ranks"])
PY
