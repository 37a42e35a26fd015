#!/bin/bash
# One GPU-box visit: parity tests, smoke, a short bench, then (only if the bench exited 0) the ncu
# launch list of the same bench command.  Everything is logged under gpurun_out/.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/gpu.txt 2>&1
timeout 900 python -m pytest tests -m gpu -q --maxfail=${MAXFAIL:-8} --timeout=300 > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/smoke.log
BENCH="python bench.py --steps ${STEPS:-10} --warmup 3"
timeout 600 $BENCH > gpurun_out/bench.log 2> gpurun_out/bench.err
rc=$?; echo "bench exit $rc"; tail -1 gpurun_out/bench.log; tail -3 gpurun_out/bench.err
if [ $rc -eq 0 ] && [ "${NCU:-1}" = "1" ]; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv \
    --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_run.log 2>&1
  echo "ncu exit $?"
fi
