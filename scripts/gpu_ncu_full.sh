#!/bin/bash
# ncu --set full captures of the hot kernels on the C2 workload (one plain run first).
set -u
mkdir -p gpurun_out
W=${NCU_WORKLOAD:-c2}
CMD="python scripts/prof_driver.py $W"
timeout 600 $CMD > gpurun_out/prof_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/prof_plain.log; exit 1; }
cat gpurun_out/prof_plain.log | tail -2
NCU="ncu --set full --clock-control none --import-source on"
# second build only: skip the first build's launches of each kernel
timeout 900 $NCU -k regex:onesweep4_kernel -s ${OS_SKIP:-27} -c 5 -f -o gpurun_out/prof_onesweep $CMD > gpurun_out/ncu_onesweep.log 2>&1; echo "onesweep rc=$?"
timeout 900 $NCU -k regex:"finish_kernel|compact_gather_kernel|query_thread_kernel|hits_kernel|pack0_codes_kernel|lcp0_kernel|hist_kernel|group_flags|writeback|encode_kernel" -s ${OTHER_SKIP:-21} -c 30 -f -o gpurun_out/prof_others $CMD > gpurun_out/ncu_others.log 2>&1; echo "others rc=$?"
ls -la gpurun_out/*.ncu-rep
