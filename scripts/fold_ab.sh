#!/bin/bash
mkdir -p gpurun_out
for f in 0 8 16; do
  echo "DLAP_FOLD_WGRAD=$f"
  DLAP_FOLD_WGRAD=$f ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/${TAG:-fa}_fold$f.csv \
    python scripts/kernel_times.py 240 10000 bf16x3 > gpurun_out/${TAG:-fa}_fold$f.log 2>&1
  python scripts/ncu_summary.py gpurun_out/${TAG:-fa}_fold$f.csv 2>/dev/null | grep "bptt\|wgrad"
  DLAP_FOLD_WGRAD=$f python bench.py --steps 20 --warmup 5 --no-legs --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('bench ms', d['ms_per_step'], 'launches', d['gpu_launches'])"
done
