#!/bin/bash
# Device annotations: parity tests (contours / layer / visualize / CLI video path), then timings of the two contour kernels.
TAG=${1:-annot}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_annotate.py tests/test_gpu_stage2.py tests/test_gpu_cli.py -m gpu -x -q --timeout 300 > gpurun_out/${TAG}_tests.log 2>&1
echo "pytest exit $?"; tail -15 gpurun_out/${TAG}_tests.log
timeout 300 python - > gpurun_out/${TAG}_timing.log 2>&1 <<'PY'
import numpy as np, torch, time, cv2
import unet_b200.synth as S
from unet_b200 import _lib
B, H, W = 64, 448, 800
c = _lib.Context(0, max_batch=B, frame_hw=(H, W), net_hw=(16, 16))
fr = torch.from_numpy(np.concatenate([S.frames_cable(16, H, W, seed=1)] * 4)).cuda()
mc = torch.from_numpy(np.stack([S.mask_cable_synth(H, W)] * B)).cuda()
mt = torch.from_numpy(np.stack([S.mask_blobs(H, W, seed=3, fill=0.02)] * B)).cuda()
for name, m in (("cable", mc), ("tape blobs", mt)):
    for _ in range(3): c.draw_contours(fr, m, (0, 255, 0))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): c.draw_contours(fr, m, (0, 255, 0))
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    mh = m[0].cpu().numpy(); img = fr[0].cpu().numpy().copy()
    t0 = time.perf_counter()
    for _ in range(20):
        cs, _ = cv2.findContours(mh, cv2.RETR_EXTERNAL, cv2.CHAIN_APPROX_SIMPLE); cv2.drawContours(img, cs, -1, (0, 255, 0), 2)
    cpu = (time.perf_counter() - t0) / 20 * 1e3
    print(f"ub_draw_contours {name}: {ms:.3f} ms per {B} frames = {ms * 1e3 / B:.2f} us/frame; cv2 on one host core {cpu * 1e3:.1f} us/frame")
PY
cat gpurun_out/${TAG}_timing.log
