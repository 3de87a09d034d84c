#!/bin/bash
# Segment-ordered floods (flood.cuh): parity tests of everything that floods, per-operator bench, contour timing, pipeline bench x2.
TAG=${1:-flood}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_annotate.py tests/test_gpu_stage2.py tests/test_gpu_configs.py tests/test_gpu_cli.py -m gpu -x -q --timeout 300 > gpurun_out/${TAG}_tests.log 2>&1
echo "pytest exit $?"; tail -3 gpurun_out/${TAG}_tests.log
timeout 300 python tools/bench_stage2.py 512 10 > gpurun_out/${TAG}_stage2.json 2> gpurun_out/${TAG}_stage2.log; echo "stage2 exit $?"; tail -10 gpurun_out/${TAG}_stage2.log | head -6
bash -c "$(sed -n '/^timeout 300 python - /,/^PY$/p' tools/gpu_annot.sh | sed "s/\${TAG}/${TAG}/g")"; cat gpurun_out/${TAG}_timing.log
for i in 1 2; do
  timeout 300 python bench.py --steps 40 --warmup 5 --no-hbm-kernels > gpurun_out/${TAG}_bench_$i.json 2> gpurun_out/${TAG}_bench_$i.err
  echo "bench run $i: $(python -c "import json,sys; d=json.loads(open('gpurun_out/${TAG}_bench_$i.json').read().strip().splitlines()[-1]); print(round(d['value']), 'e2e', round(d['e2e']['value']), d['clocks']['sm_mhz'])")"
done
