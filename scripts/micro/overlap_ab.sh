#!/bin/bash
# A/B of functional.overlap_macro on one GPU box: the bit-identity test, then the bench with the LSTM
# recurrence beside the per-cell forward kernel and one after the other.
set -u
mkdir -p gpurun_out
timeout -k 10 900 python -m pytest tests -m gpu -q -x --timeout=300 > gpurun_out/ab_pytest.log 2>&1
echo "pytest exit $?"; tail -3 gpurun_out/ab_pytest.log
for ov in 1 0 1 0; do
  DLAP_OVERLAP_MACRO=$ov timeout 300 python bench.py --steps 300 --warmup 3 --no-cpu > gpurun_out/ab_bench_$ov.log 2> gpurun_out/ab_bench_$ov.err
  echo "overlap=$ov exit $?"; python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/ab_bench_$ov.log").read().strip().splitlines()[-1])
    print({k: d[k] for k in ("value", "ms_per_step", "gpu_launches")}, d["e2e"]["value"], d["roofline"]["achieved"])
except Exception as e:
    print("no line", e); print(open("gpurun_out/ab_bench_$ov.err").read()[-1500:])
PY
done
