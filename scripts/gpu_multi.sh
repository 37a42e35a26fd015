#!/bin/bash
# Multi-GPU visit on one box with G GPUs: NCCL parity tests, bench at every N in NS, per-entry-point breakdown at G.
set -u
TAG=${TAG:-multi}; G=${G:-8}; NS=${NS:-"2 4 8"}
mkdir -p gpurun_out
if [ "${TESTS:-1}" = "1" ]; then
  timeout 240 python -m pytest tests/test_sharding_gpu.py -m gpu -q -x -k nccl > gpurun_out/${TAG}_nccl_tests.log 2>&1; echo "nccl tests exit $?"; tail -2 gpurun_out/${TAG}_nccl_tests.log
fi
for n in $NS; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
    bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/${TAG}_bench$n.json 2> gpurun_out/${TAG}_bench$n.err
  echo "bench N=$n exit $?"
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/${TAG}_bench$n.json").read().strip().split("\n")[-1])
    print("  value %.3e  ms %.4f  e2e %.3e  scaled %s  check %s" % (d["value"], d["ms_per_step"], d["e2e"]["value"],
          {k: d.get("scaled_panel", {}).get(k) for k in ("value", "ms_per_step")}, d.get("sharded_check", {}).get("loss_rel")))
except Exception as e:
    print("  no line:", e)
PY
done
[ "${EXTRAS:-1}" = "1" ] || exit 0
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29600 scripts/step_breakdown.py > gpurun_out/${TAG}_breakdown$G.json 2> gpurun_out/${TAG}_breakdown$G.err; echo "breakdown exit $?"; tail -1 gpurun_out/${TAG}_breakdown$G.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29601 scripts/ensemble_gpu.py > gpurun_out/${TAG}_ensemble$G.json 2> gpurun_out/${TAG}_ensemble$G.err; echo "ensemble exit $?"; tail -1 gpurun_out/${TAG}_ensemble$G.json
