#!/bin/bash
# kernel change check: build parity grid + config-size parity, then the N=1 bench
set -u
mkdir -p gpurun_out
( time python -m pytest tests/test_gpu_build_parity.py tests/test_gpu_config_parity.py tests/test_gpu_sharded.py -x -q -m gpu ) > gpurun_out/r2_kern_tests.log 2>&1; echo "tests rc=$?"; tail -6 gpurun_out/r2_kern_tests.log
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench_n1_kern.json 2> gpurun_out/r2_bench_n1_kern.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.load(open("gpurun_out/r2_bench_n1_kern.json"))
print(d["value"], d["build_ms_per_step"], d["phases_ms_last_step"], d.get("parity"))
PY
