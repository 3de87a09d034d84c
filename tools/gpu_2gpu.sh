#!/bin/bash
# 2-GPU visit: torchrun bench (contract launch line) for both arms + the drop-in CLI under torchrun.
mkdir -p gpurun_out
python - <<'PY'
import torch, sys
sys.path.insert(0, '.')
import unet_b200.synth as S
torch.save({'model': S.make_state_dict(0)}, '/tmp/ckpt.pth')
PY
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r02g_n2.json 2> gpurun_out/r02g_n2.err; echo "bench N=2 exit $?"; wc -l gpurun_out/r02g_n2.json; cut -c1-250 gpurun_out/r02g_n2.json
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > gpurun_out/r02g_n2_ref.json 2> gpurun_out/r02g_n2_ref.err; echo "ref N=2 exit $?"; wc -l gpurun_out/r02g_n2_ref.json; cut -c1-200 gpurun_out/r02g_n2_ref.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 -m unet_b200.cli_two_stage --synthetic 200 --model /tmp/ckpt.pth --output /tmp/out2 --no-video-out --batch 32 > gpurun_out/r02g_cli_2gpu.log 2>&1; echo "cli 2gpu exit $?"; tail -6 gpurun_out/r02g_cli_2gpu.log
timeout 300 python -m unet_b200.cli_two_stage --synthetic 200 --model /tmp/ckpt.pth --output /tmp/out1 --no-video-out --batch 32 > gpurun_out/r02g_cli_1gpu.log 2>&1; echo "cli 1gpu exit $?"; tail -6 gpurun_out/r02g_cli_1gpu.log
