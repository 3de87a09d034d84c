"""Times the stage-2 entry points one by one through the C-ABI (CUDA events, gray planes resident in
HBM) and prints achieved GB/s on each call's own compulsory bytes (1 B/px per u8 plane
read or written).  B = 512 frames (183 MB per plane, larger than the 126 MB L2) unless given.
Usage: python tools/bench_stage2.py [frames] [reps]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import unet_b200.synth as S
from unet_b200 import _lib

B = int(sys.argv[1]) if len(sys.argv) > 1 else 512
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
H, W = 448, 800
P = H * W
torch.cuda.set_device(0)
ctx = _lib.Context(0, max_batch=B, frame_hw=(H, W), net_hw=(16, 16))   # stage 2 only: keep the network's activation planes tiny
base = np.concatenate([S.frames_cable(12, H, W, seed=1), S.frames_noise(4, H, W, seed=2)])
gray16 = np.stack([(f.astype(np.uint32) @ np.array([3735, 19235, 9798], np.uint32) + (1 << 14) >> 15).astype(np.uint8) for f in base])
gray = torch.from_numpy(np.concatenate([gray16] * (B // 16))).cuda()
mask = torch.from_numpy(np.stack([S.mask_cable_synth(H, W)] * B)).cuda()
peaks = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json"))) if os.path.exists(
    os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) else {"hbm_gbs": 6549.8}


def timed(fn):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(torch.cuda.current_stream())
    for _ in range(reps):
        fn()
    e1.record(torch.cuda.current_stream())
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


frames3 = torch.from_numpy(np.concatenate([base] * (B // 16))).cuda()
mask_t = torch.from_numpy(np.stack([S.mask_blobs(H, W, seed=3, fill=0.02)] * B)).cuda()
mask_b = torch.from_numpy(np.stack([S.mask_blobs(H, W, seed=4, fill=0.004)] * B)).cuda()
roi = (218, 0, 421, 448)
rows = [
    ("overlay blend (4 x addWeighted, ub_overlay)", lambda: ctx.overlay(frames3, mask, mask_t, mask_b, roi), 9 * P),
    ("gray (ub_preprocess, gray only)", lambda: ctx.preprocess(frames3, want_x=False, want_gray=True), 4 * P),
    ("gaussian5 (k_edge_fused<2>)", lambda: ctx.gaussian5(gray), 2 * P),
    ("blur_canny, hysteresis included", lambda: ctx.blur_canny(gray, 50, 150), 2 * P),
    ("burr stage (edges + band + morphology + CCL + filter)", lambda: ctx.burr(gray, mask), 3 * P),
    ("morph CLOSE 3x3", lambda: ctx.morph("close", 3, mask), 2 * P),
    ("morph dilate 8x8", lambda: ctx.morph("dilate", 8, mask), 2 * P),
]
out = {"frames": B, "hw": [H, W], "hbm_peak_gbs": peaks["hbm_gbs"], "fused": os.environ.get("UB_EDGE_FUSED", "1"), "ops": []}
for name, fn, nbytes in rows:
    ms = timed(fn)
    gbs = nbytes * B / ms / 1e6
    out["ops"].append({"op": name, "ms": round(ms, 4), "us_per_frame": round(ms * 1e3 / B, 3), "algorithmic_bytes_per_frame": nbytes,
                       "achieved_gbs": round(gbs, 1), "frac_of_hbm_peak": round(gbs / peaks["hbm_gbs"], 4)})
    print(f"{name:55s} {ms:8.3f} ms  {ms * 1e3 / B:7.2f} us/frame  {gbs:8.1f} GB/s  {gbs / peaks['hbm_gbs']:.3f} of HBM peak", file=sys.stderr)
print(json.dumps(out))
