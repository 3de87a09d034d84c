#!/bin/bash
# enhanced workload bench, CLI as a module, UB_TC_PAIR A/B
mkdir -p gpurun_out
timeout 400 python bench.py --workload enhanced --steps 6 --warmup 3 --cpu-frames 4 > gpurun_out/misc_enh_bench.json 2> gpurun_out/misc_enh_bench.err
echo "enhanced bench exit $?"; cat gpurun_out/misc_enh_bench.json; tail -2 gpurun_out/misc_enh_bench.err
python - <<'PY'
import torch, sys
sys.path.insert(0, '.')
import unet_b200.synth as S
torch.save({'model': S.make_state_dict(0)}, '/tmp/ckpt.pth')
PY
timeout 300 python -m unet_b200.cli_two_stage --synthetic 300 --model /tmp/ckpt.pth --output /tmp/out1 --no-video-out > gpurun_out/misc_cli1.log 2>&1; echo "cli1 exit $?"; tail -6 gpurun_out/misc_cli1.log
timeout 300 python -m unet_b200.cli_two_stage --synthetic 40 --model /tmp/ckpt.pth --output /tmp/out2 --batch 16 > gpurun_out/misc_cli2.log 2>&1; echo "cli2 exit $?"; tail -3 gpurun_out/misc_cli2.log; ls -la /tmp/out2
timeout 300 python -m unet_b200.cli_enhanced --synthetic 32 --synthetic-size 1920x1080 --model /tmp/ckpt.pth --output /tmp/out3 --enhance --no-video-out > gpurun_out/misc_cli3.log 2>&1; echo "cli3 exit $?"; tail -4 gpurun_out/misc_cli3.log
bash tools/gpu_sweep.sh pair UB_TC_PAIR 0 1
