"""ub_draw_contours on 64 frames of 800x448 (cable mask, blob mask): CUDA-event time per frame, cv2 on one host core beside it."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cv2
import numpy as np
import torch

import unet_b200.synth as S
from unet_b200 import _lib

B, H, W = 64, 448, 800
c = _lib.Context(0, max_batch=B, frame_hw=(H, W), net_hw=(16, 16))
fr = torch.from_numpy(np.concatenate([S.frames_cable(16, H, W, seed=1)] * 4)).cuda()
masks = (("cable", torch.from_numpy(np.stack([S.mask_cable_synth(H, W)] * B)).cuda()),
         ("tape blobs", torch.from_numpy(np.stack([S.mask_blobs(H, W, seed=3, fill=0.02)] * B)).cuda()))
for name, m in masks:
    for _ in range(3):
        c.draw_contours(fr, m, (0, 255, 0))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        c.draw_contours(fr, m, (0, 255, 0))
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    mh, img = m[0].cpu().numpy(), fr[0].cpu().numpy().copy()
    t0 = time.perf_counter()
    for _ in range(20):
        cs, _ = cv2.findContours(mh, cv2.RETR_EXTERNAL, cv2.CHAIN_APPROX_SIMPLE)
        cv2.drawContours(img, cs, -1, (0, 255, 0), 2)
    cpu = (time.perf_counter() - t0) / 20 * 1e3
    print(f"ub_draw_contours {name}: {ms:.3f} ms per {B} frames = {ms * 1e3 / B:.2f} us/frame; cv2 on one host core {cpu * 1e3:.1f} us/frame")
