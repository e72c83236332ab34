#!/bin/bash
# end of round: whole GPU suite, smoke, the N=1 bench line and the reference arm
set -u
mkdir -p gpurun_out
( time python -m pytest tests -x -q -m gpu ) > gpurun_out/r2_pytest_gpu_final.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/r2_pytest_gpu_final.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke_final.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2_smoke_final.log
python bench.py > gpurun_out/r2_bench_n1_final.json 2> gpurun_out/r2_bench_n1_final.err; echo "bench rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2_bench_ref_final.json 2> gpurun_out/r2_bench_ref_final.err; echo "ref rc=$?"
python - <<'PY'
import json
d = json.load(open("gpurun_out/r2_bench_n1_final.json"))
print(d["value"], d["build_ms_per_step"], d["e2e"]["value"], d["phases_ms_last_step"], d["roofline"]["frac"], d["roofline_query"]["frac"], d["parity_checked"], d["clocks"])
r = json.load(open("gpurun_out/r2_bench_ref_final.json"))
print("ref", r["value"], r.get("cpu_baseline", {}).get("cores"))
PY
