#!/bin/bash
# N GPUs: cost-weighted splitters (gamma sweep); first the emulated-rank parity tests on GPU 0
set -u
N=${1:-8}
mkdir -p gpurun_out
( CUDA_VISIBLE_DEVICES=0 timeout 300 python -m pytest tests/test_gpu_multi_native.py -x -q -m gpu 2>&1 | tail -3 )
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
i=0
for g in 0 110 140 170; do
  i=$((i+1))
  SA_BENCH_DEBUG="split_gamma_x100=$g" timeout 300 $TR --master-port 2955$i bench.py --gpus $N --steps 4 --warmup 3 --no-cpu-baseline \
     > gpurun_out/r2_n${N}_gamma_$g.json 2> gpurun_out/r2_n${N}_gamma_$g.err; echo "gamma=$g rc=$?"
  python - <<PY
import json
d = json.load(open("gpurun_out/r2_n${N}_gamma_$g.json"))
m = d["multi_gpu"]
print("gamma $g build", d["build_ms_per_step"], "e2e", d["e2e"]["build_ms"], "imbalance", m["imbalance"], m["phases_ms_max_over_ranks"])
print("  sort+refine", m["phases_ms_by_rank"]["sort+refine"], "finish", m["phases_ms_by_rank"]["finish+replicate"][:2])
PY
done
