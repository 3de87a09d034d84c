#!/bin/bash
# compute-sanitizer over the stage-2 parity tests that exercise the new kernels, then a short bench (clock sampler check).
TAG=${1:-san}
mkdir -p gpurun_out
for tool in memcheck racecheck; do
  timeout 900 compute-sanitizer --tool $tool --error-exitcode 99 --print-limit 20 \
    python -m pytest tests/test_gpu_stage2.py -m gpu -x -q --timeout 800 -k "gauss or overlay or rulebased or visualize or burr_stage_batched" \
    > gpurun_out/${TAG}_$tool.log 2>&1
  echo "$tool exit $?"; grep -E "ERROR SUMMARY|passed|failed|Error" gpurun_out/${TAG}_$tool.log | head -8
done
timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-hbm-kernels > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"
python -c "import json; d=json.load(open('gpurun_out/${TAG}_bench.json')); print(d['value'], d['clocks'])"
