#!/bin/bash
# Conv metrics pass (fresh profiles/rNN_conv_ncu.json for bench.py's roofline.traffic) + the annotate tests + a bench line.
TAG=${1:-convm}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_annotate.py -m gpu -x -q --timeout 300 > gpurun_out/${TAG}_annot_tests.log 2>&1
echo "pytest exit $?"; tail -2 gpurun_out/${TAG}_annot_tests.log
bash tools/gpu_conv_metrics.sh ${TAG}
