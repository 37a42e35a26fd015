#!/bin/bash
# block size of the forward kernel's block-cyclic tile order (2^sh tiles): the forward pieces, then the graphed step
for sh in 0 1 2 3; do
  echo "sh=$sh"; DLAP_CYC_SH=$sh timeout 200 python scripts/micro/fwd_overlap_probe.py 2>&1 | tail -1
  DLAP_CYC_SH=$sh timeout 300 python bench.py --steps 300 --warmup 3 --no-cpu 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['e2e']['value'])"
done
