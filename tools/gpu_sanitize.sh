#!/bin/bash
# compute-sanitizer over the tcgen05 / TMA / mbarrier kernels (hand-rolled barrier phases, cluster multicast, cross-warp
# shared-memory exchanges) and the stage-2 kernels.  Every run under its own timeout; the ERROR SUMMARY lines are what
# profiles/r02_sanitize.md quotes.
TAG=${1:-san}
mkdir -p gpurun_out
run() {  # name tool timeout pytest-args...
  local name=$1 tool=$2 to=$3; shift 3
  timeout $to compute-sanitizer --tool $tool --error-exitcode 99 --print-limit 10 \
    python -m pytest "$@" -m gpu -x -q --timeout $to -p no:cacheprovider > gpurun_out/${TAG}_${name}_${tool}.log 2>&1
  echo "== $name $tool exit $?"; grep -E "ERROR SUMMARY|passed|failed|Error:|Race reported|hazard" gpurun_out/${TAG}_${name}_${tool}.log | head -6
}
# layers 0/1 (conv_rows, fold), 5 (conv_tc N = 128, CTA pair), 10 (conv_tc N = 256, two sources), 14/16 (fused upsample), 17 (head)
NET='tests/test_gpu_net.py -k "test_conv_layer and tcgen05 and not aligned and (0-t or 1-t or 5-t or 10-t or 14-t or 16-t or 17-t) or fused_block or fused_pool"'
for tool in memcheck racecheck synccheck; do
  eval run net $tool 420 $NET
done
eval run ups memcheck 300 'tests/test_gpu_net.py -k "fused_upsample"'
eval run ups racecheck 300 'tests/test_gpu_net.py -k "fused_upsample and 16-hw0"'
run s2 memcheck 420 tests/test_gpu_stage2.py -k "gauss or overlay or rulebased or burr_stage_batched or morphology or ccl"
run s2 racecheck 420 tests/test_gpu_stage2.py -k "gauss or burr_stage_batched or ccl"
