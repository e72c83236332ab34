#!/bin/bash
# N GPUs (default 4): where finish + replicate goes (timeline of finished / pushed chunks); push kernel and chunk count A/B
set -u
N=${1:-4}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
i=0
for cfg in "replicate_ctas=64" "replicate_ctas=0" "replicate_ctas=0,finish_chunks=16"; do
  i=$((i+1))
  SA_BENCH_DEBUG="timeline=1,$cfg" timeout 300 $TR --master-port 2954$i bench.py --gpus $N --steps 3 --warmup 3 --no-cpu-baseline \
     > gpurun_out/r2_n${N}_diag_$i.json 2> gpurun_out/r2_n${N}_diag_$i.err; echo "cfg=$cfg rc=$?"
  grep "timeline rank 0" gpurun_out/r2_n${N}_diag_$i.err | tail -1
  python - <<PY
import json
d = json.load(open("gpurun_out/r2_n${N}_diag_$i.json"))
print("$cfg build", d["build_ms_per_step"], "e2e", d["e2e"]["build_ms"], d["multi_gpu"]["phases_ms_max_over_ranks"])
print(d["multi_gpu"]["phases_ms_by_rank"])
PY
done
