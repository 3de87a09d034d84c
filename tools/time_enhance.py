"""Time ub_enhance (and its steps) on the device: python tools/time_enhance.py [batch]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import unet_b200.synth as S
from unet_b200 import _lib

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
c = _lib.Context(0, max_batch=B, frame_hw=(448, 800), net_hw=(512, 512))
fr = torch.from_numpy(S.frames_cable(B, 448, 800, seed=0)).cuda()


def t(fn, n=3):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n


lab = c.cvt_lab(fr, 1)
l, ab = lab[..., 0].contiguous(), lab[..., 1:].contiguous()
print(f"batch {B}")
print(f"cvt_lab     {t(lambda: c.cvt_lab(fr, 0)) / B:8.4f} ms/frame")
print(f"clahe       {t(lambda: c.clahe(l)) / B:8.4f} ms/frame")
print(f"nlm L       {t(lambda: c.nlm(l)) / B:8.4f} ms/frame")
print(f"nlm ab      {t(lambda: c.nlm(ab)) / B:8.4f} ms/frame")
print(f"sharpen     {t(lambda: c.sharpen3(fr)) / B:8.4f} ms/frame")
print(f"enhance     {t(lambda: c.enhance(fr)) / B:8.4f} ms/frame  (reference CPU: ~510 ms/frame)")
