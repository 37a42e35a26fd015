#!/bin/bash
# e2e leg of the bench: LSTM beside cell_fwd or not, hardware queue count
for cfg in "1 8" "1 32" "0 8" "0 32"; do set -- $cfg
  CUDA_DEVICE_MAX_CONNECTIONS=$2 DLAP_OVERLAP_MACRO=$1 timeout 300 python bench.py --steps 100 --warmup 3 --no-cpu 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('overlap $1 connections $2', d['ms_per_step'], d['e2e']['ms_per_step'], d['e2e']['value'])"
done
