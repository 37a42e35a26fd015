#!/bin/bash
# Forward kernels timed separately: the macro LSTM as its own launch (DLAP_FUSE_MACRO=0) against the fused launch.
mkdir -p gpurun_out
for f in 0 1; do
  DLAP_FUSE_MACRO=$f ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/${TAG:-fs}_fuse$f.csv \
    python scripts/kernel_times.py 240 10000 bf16x3 > gpurun_out/${TAG:-fs}_fuse$f.log 2>&1
  python scripts/ncu_summary.py gpurun_out/${TAG:-fs}_fuse$f.csv 2>/dev/null | tail -12
done
