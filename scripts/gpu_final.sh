#!/bin/bash
# Final check of a round: GPU parity tests, smoke, default bench line, reference arm, then the ncu
# launch list (+ DRAM bytes) of the CSV path (plain run first, as B200_PROFILING.md requires).
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 500 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"
CMD="python scripts/prof_csv.py --csv-mb 256"
timeout 300 $CMD > gpurun_out/prof_csv_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 2000 --csv \
    --log-file gpurun_out/launches_csv.csv $CMD > gpurun_out/ncu_csv.log 2>&1
echo "ncu csv rc=$?"; tail -1 gpurun_out/prof_csv_plain.log
cut -c1-400 gpurun_out/bench.json; cut -c1-300 gpurun_out/bench_ref.json
