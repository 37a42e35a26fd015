#!/bin/bash
# One GPU-box visit: parity tests, smoke, the bench, then (only if the bench exited 0) the ncu launch
# list of the same bench command and one --set full capture of the dominant kernel.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/gpu.txt 2>&1
timeout -k 10 900 python -m pytest tests -m gpu -q --maxfail=${MAXFAIL:-8} --timeout=300 > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?"; grep -E "passed|failed" gpurun_out/pytest_gpu.log | tail -2
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -1 gpurun_out/smoke.log
timeout 600 python bench.py --steps ${STEPS:-100} --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err
rc=$?; echo "bench exit $rc"; tail -1 gpurun_out/bench.log; tail -3 gpurun_out/bench.err
if [ $rc -eq 0 ] && [ "${NCU:-1}" = "1" ]; then
  BENCH="python bench.py --steps 2 --warmup 3 --no-cpu"
  timeout 300 $BENCH > gpurun_out/ncu_plain.log 2>&1 && \
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/launches.csv $BENCH > gpurun_out/ncu_run.log 2>&1
  echo "ncu launches exit $?"
  timeout 300 $BENCH > gpurun_out/ncu_plain.log 2>&1 && \
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:cell_bwd_tc -s 3 -c 1 \
    -o gpurun_out/dominant -f $BENCH > gpurun_out/ncu_full.log 2>&1
  echo "ncu full exit $?"
fi
