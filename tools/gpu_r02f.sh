#!/bin/bash
TAG=${1:-r02f}
mkdir -p gpurun_out
timeout 60 ./build/ubench_shfl > gpurun_out/${TAG}_ubench_shfl.log 2>&1; cat gpurun_out/${TAG}_ubench_shfl.log
timeout 600 python -m pytest tests/test_gpu_net.py tests/test_gpu_configs.py -m gpu -x -q --timeout 120 > gpurun_out/${TAG}_net_tests.log 2>&1
echo "net tests exit $?"; tail -4 gpurun_out/${TAG}_net_tests.log
timeout 200 python bench.py --quick --no-cpu-baseline --no-hbm-kernels --steps 60 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
python - <<PY
import json
d = json.load(open("gpurun_out/${TAG}_bench.json"))
print("frames/s", round(d["value"], 1), "ms/step", round(d["ms_per_step"], 3), "conv ms", round(d["roofline"]["leg"]["conv_ms_per_step"], 3), d["clocks"]["sm_mhz"], "launches", d["gpu_launches"])
PY
bash tools/gpu_conv_metrics.sh ${TAG}
bash tools/gpu_sanitize.sh ${TAG}
