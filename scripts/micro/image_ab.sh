#!/bin/bash
# X images on / off: parity tests of the image path, then per-kernel times (ncu launch list) both ways.
mkdir -p gpurun_out
TAG=${TAG:-img}
timeout 600 python -m pytest tests/test_x_image_gpu.py -m gpu -q -x > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/${TAG}_pytest.log
for f in 0 1; do
  DLAP_X_IMAGE=$f ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/${TAG}_image$f.csv \
    python scripts/kernel_times.py 240 10000 bf16x3 > gpurun_out/${TAG}_image$f.log 2>&1
  echo "DLAP_X_IMAGE=$f"; python scripts/ncu_summary.py gpurun_out/${TAG}_image$f.csv 2>/dev/null | grep dlap
  DLAP_X_IMAGE=$f python scripts/kernel_times.py 240 10000 bf16x3
done
