#!/bin/bash
mkdir -p gpurun_out
TAG=${TAG:-fp}
python scripts/bwd_timeline.py > gpurun_out/${TAG}_timeline_fused.txt 2>&1; grep "FWD\|BWD gather\|tile period\|MMA:" gpurun_out/${TAG}_timeline_fused.txt
DLAP_FUSE_MACRO=0 python scripts/bwd_timeline.py > gpurun_out/${TAG}_timeline_unfused.txt 2>&1; grep "FWD" gpurun_out/${TAG}_timeline_unfused.txt
DLAP_FUSE_MACRO=0 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/${TAG}_unfused.csv \
    python scripts/kernel_times.py 240 10000 bf16x3 > gpurun_out/${TAG}_unfused.log 2>&1
python scripts/ncu_summary.py gpurun_out/${TAG}_unfused.csv 2>/dev/null | grep dlap
