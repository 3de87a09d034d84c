#!/bin/bash
# 8-GPU visit: both bench arms under the contract's torchrun line.   Usage: bash tools/gpu_8gpu.sh TAG   (gpurun --gpus 8)
TAG=${1:-r02x}
mkdir -p gpurun_out
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/${TAG}_n8.json 2> gpurun_out/${TAG}_n8.err; echo "bench N=8 exit $?"; wc -l gpurun_out/${TAG}_n8.json; cut -c1-260 gpurun_out/${TAG}_n8.json; tail -3 gpurun_out/${TAG}_n8.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29522 bench.py --impl reference --gpus 8 --steps 3 --warmup 1 > gpurun_out/${TAG}_n8_ref.json 2> gpurun_out/${TAG}_n8_ref.err; echo "ref N=8 exit $?"; cut -c1-200 gpurun_out/${TAG}_n8_ref.json
