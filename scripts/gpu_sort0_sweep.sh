set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_csv_columns.py tests/test_gpu_sharded.py tests/test_gpu_csv_scan.py tests/test_gpu_dropin.py -m gpu -q > gpurun_out/pytest_csv.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_csv.log
for c in 12 11 10 9 8 6; do
  SA_BENCH_DEBUG=sort0_chars=$c timeout 300 python bench.py --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/bench_s0_$c.json 2> gpurun_out/bench_s0_$c.err
  python - <<PY
import json
b=json.load(open("gpurun_out/bench_s0_$c.json"))
print($c, b["value"], b["build_ms_per_step"], b["phases_ms_last_step"], b["refine_rounds"], b["roofline"]["launches_per_step"], b["roofline"]["achieved"])
PY
done
