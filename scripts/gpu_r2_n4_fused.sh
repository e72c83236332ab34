#!/bin/bash
# N GPUs (default 4): finish_kernel storing into every replica itself vs finishing locally + replicate_kernel
set -u
N=${1:-4}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
i=0
for cfg in "replicate_fused=1" "replicate_fused=0"; do
  i=$((i+1))
  extra=""; [ $i -eq 1 ] && extra="--verify-single"
  SA_BENCH_DEBUG="$cfg" timeout 300 $TR --master-port 2956$i bench.py --gpus $N --steps 3 --warmup 3 --no-cpu-baseline $extra \
     > gpurun_out/r2_n${N}_fused_$i.json 2> gpurun_out/r2_n${N}_fused_$i.err; echo "cfg=$cfg rc=$?"
  python - <<PY
import json
d = json.load(open("gpurun_out/r2_n${N}_fused_$i.json"))
print("$cfg build", d["build_ms_per_step"], "e2e", d["e2e"]["build_ms"], d["multi_gpu"]["phases_ms_max_over_ranks"], d.get("verify"))
PY
done
