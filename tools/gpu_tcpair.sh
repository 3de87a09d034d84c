#!/bin/bash
# A/B of k_conv_tc pair modes: network parity tests with the default (cta_group::2), then quick benches with UB_TC_PAIR=1 / 2.
TAG=${1:-r02c}
mkdir -p gpurun_out
UB_TC_PAIR=3 timeout 600 python -m pytest tests/test_gpu_net.py -m gpu -x -q --timeout 120 -k "conv_layer or forward_vs" > gpurun_out/${TAG}_net_tests.log 2>&1
echo "pytest exit $?"; tail -5 gpurun_out/${TAG}_net_tests.log
for mode in 1 2 3 4; do
  UB_TC_PAIR=$mode timeout 200 python bench.py --quick --no-cpu-baseline --no-hbm-kernels --steps 60 > gpurun_out/${TAG}_bench_pair${mode}.json 2> gpurun_out/${TAG}_bench_pair${mode}.err
  python - <<PY
import json
d = json.load(open("gpurun_out/${TAG}_bench_pair${mode}.json"))
print("UB_TC_PAIR=${mode}: frames/s", round(d["value"], 1), "ms/step", round(d["ms_per_step"], 3), "conv ms", round(d["roofline"]["leg"]["conv_ms_per_step"], 3), d["clocks"]["sm_mhz"])
PY
done
UB_TC_PAIR=3 bash tools/gpu_conv_metrics.sh ${TAG}
