#!/bin/bash
# BASELINE configs 4 and 5 on N GPUs (torchrun per the bench contract) + the new module-configuration test on GPU 0.
N=${1:-8}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_net.py -m gpu -x -q --timeout 120 -k "other_configurations or other_class" 2>&1 | tail -3
for wl in net512 enhanced; do
  for n in $N 1; do
    if [ "$n" = "1" ]; then
      timeout 400 python bench.py --workload $wl --steps 8 --warmup 3 --no-cpu-baseline --no-hbm-kernels > gpurun_out/r01r_${wl}_n1.json 2> gpurun_out/r01r_${wl}_n1.err
    else
      timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29519 \
        bench.py --gpus $n --workload $wl --steps 8 --warmup 3 --no-cpu-baseline --no-hbm-kernels > gpurun_out/r01r_${wl}_n$n.json 2> gpurun_out/r01r_${wl}_n$n.err
    fi
    echo "$wl N=$n exit $?"; python -c "
import json; d=json.load(open('gpurun_out/r01r_${wl}_n$n.json')); print(d['config']['workload'], d['n_gpus'], round(d['value'],1), round(d['e2e']['value'],1), round(d['ms_per_step'],2), d['clocks'])" || tail -5 gpurun_out/r01r_${wl}_n$n.err
  done
done
