#!/bin/bash
# per-kernel times (ncu launch list) + eager step time of the paper-shape step
mkdir -p gpurun_out
TAG=${TAG:-qt}
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/${TAG}_launches.csv \
    python scripts/kernel_times.py 240 10000 bf16x3 > gpurun_out/${TAG}_kt.log 2>&1
python scripts/ncu_summary.py gpurun_out/${TAG}_launches.csv 2>/dev/null | grep dlap
python bench.py --steps 20 --warmup 5 --no-legs --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('bench ms', d['ms_per_step'], 'value', d['value'], 'launches', d['gpu_launches'], 'bwd ms', d['roofline']['kernel_ms'])"
