#!/bin/bash
# Fused conv0_0 block: parity tests (each under its own timeout: a protocol bug in a persistent kernel hangs), then A/B bench.
TAG=${1:-r02e}
mkdir -p gpurun_out
timeout 180 python -m pytest tests/test_gpu_net.py -m gpu -x -q --timeout 60 -k "fused_block or forward_on_frames or pipeline_vs_oracle" > gpurun_out/${TAG}_block_tests.log 2>&1
rc=$?; echo "block tests exit $rc"; tail -15 gpurun_out/${TAG}_block_tests.log
if [ $rc -ne 0 ]; then exit 0; fi
timeout 600 python -m pytest tests/test_gpu_net.py tests/test_gpu_configs.py -m gpu -x -q --timeout 120 > gpurun_out/${TAG}_net_tests.log 2>&1
echo "net tests exit $?"; tail -4 gpurun_out/${TAG}_net_tests.log
for fb in 0 1 0 1; do
  UB_FUSE_BLOCK0=$fb timeout 200 python bench.py --quick --no-cpu-baseline --no-hbm-kernels --steps 60 > gpurun_out/${TAG}_bench_fb${fb}.json 2> gpurun_out/${TAG}_bench_fb${fb}.err
  python - <<PY
import json
d = json.load(open("gpurun_out/${TAG}_bench_fb${fb}.json"))
print("UB_FUSE_BLOCK0=${fb}: frames/s", round(d["value"], 1), "ms/step", round(d["ms_per_step"], 3), "conv ms", round(d["roofline"]["leg"]["conv_ms_per_step"], 3), d["clocks"]["sm_mhz"], "launches", d["gpu_launches"])
PY
done
bash tools/gpu_conv_metrics.sh ${TAG}
