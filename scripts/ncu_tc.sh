#!/bin/bash
# ncu --set full of the two tensor-core per-cell kernels (one launch each), after a plain run.
set -u
mkdir -p gpurun_out
CMD="python scripts/kernel_times.py ${T:-240} ${N:-10000} bf16x3"
timeout 300 $CMD > gpurun_out/ncu_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:cell_.*_tc -s 6 -c 2 \
   -o gpurun_out/tc_prof -f $CMD > gpurun_out/ncu_tc.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/ncu_tc.log
