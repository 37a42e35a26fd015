#!/bin/bash
# One GPU-box visit: parity tests, smoke, the bench (both arms).  TAG names the output files under gpurun_out/.
set -u
TAG=${TAG:-visit}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${TAG}_gpu.txt 2>&1
timeout -k 10 ${TEST_TIMEOUT:-1200} python -m pytest tests -m gpu -q -s --maxfail=${MAXFAIL:-10} ${PYTEST_ARGS:-} > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest exit $?"; grep -E "passed|failed|error" gpurun_out/${TAG}_pytest.log | tail -3
timeout 300 python __graft_entry__.py smoke > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -4 gpurun_out/${TAG}_smoke.log
timeout 900 python bench.py --steps ${STEPS:-20} --warmup 5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench exit $?"; tail -c 1500 gpurun_out/${TAG}_bench.json; tail -5 gpurun_out/${TAG}_bench.err
if [ "${REF:-0}" = "1" ]; then
  timeout 600 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/${TAG}_ref.json 2> gpurun_out/${TAG}_ref.err
  echo "ref exit $?"; tail -c 400 gpurun_out/${TAG}_ref.json
fi
if [ "${NCU:-0}" = "1" ]; then
  # launch list of the bench command (per-launch durations, cold cache, serialised) and one full capture of the dominant kernel
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv \
    python bench.py --steps 20 --warmup 5 --no-legs --no-cpu > gpurun_out/${TAG}_ncu_bench.log 2>&1
  echo "ncu launches exit $?"; python scripts/ncu_summary.py gpurun_out/${TAG}_launches.csv | grep dlap
  TAG=${TAG}_bwd KERN=cell_bwd_tc SKIP=3 bash scripts/ncu_full.sh | grep -v "^  smsp__average" | head -40
fi
