"""Runs a few steps of the bench workload (two-stage pipeline, 800x448, batch 64) for ncu.
Usage: python tools/profile_step.py [steps] [batch]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
import unet_b200.synth as S
from unet_b200.pipeline import InspectionPipeline

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
batch = int(sys.argv[2]) if len(sys.argv) > 2 else 64
torch.cuda.set_device(0)
pipe = InspectionPipeline(S.make_state_dict(0), batch=batch)
frames = torch.from_numpy(bench.make_frames(batch, 0)).cuda()
for _ in range(steps):
    stats, _ = pipe.run(frames)
torch.cuda.synchronize()
print("done", stats[:, 1:].sum(0).tolist(), "launches", pipe.ctx.launch_count())
