#!/bin/bash
# Scaling visit on an N-GPU box:  bash tools/gpu_scale.sh TAG N1 N2 ...   (torchrun per the bench contract)
TAG=$1; shift
mkdir -p gpurun_out
for n in "$@"; do
  if [ "$n" = "1" ]; then
    timeout 300 python bench.py --gpus 1 --steps 30 --warmup 5 --no-cpu-baseline --no-hbm-kernels > gpurun_out/${TAG}_n1.json 2> gpurun_out/${TAG}_n1.err
  else
    timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 \
      bench.py --gpus $n --steps 30 --warmup 5 --no-cpu-baseline --no-hbm-kernels > gpurun_out/${TAG}_n$n.json 2> gpurun_out/${TAG}_n$n.err
  fi
  echo "N=$n exit $?"; tail -1 gpurun_out/${TAG}_n$n.json | cut -c1-400
done
