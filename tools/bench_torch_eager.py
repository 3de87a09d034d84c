"""Same-box library comparison: the network of src/models/unetpp.py written with stock torch.nn layers
(cuDNN / ATen, what the reference itself runs on a GPU) against ub_forward / the net part of ub_pipeline, same
synthetic checkpoint, batch 64 at 512x512, CUDA events.  Measurement tool only -- nothing here is on the product path.
Usage: python tools/bench_torch_eager.py [batch] [reps]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.nn as nn
import torch.nn.functional as F

import unet_b200.synth as S
from unet_b200 import NestedUNet

W = (32, 64, 128, 256, 512)


class Block(nn.Module):
    def __init__(self, cin, cout):
        super().__init__()
        self.conv1, self.bn1 = nn.Conv2d(cin, cout, 3, padding=1), nn.BatchNorm2d(cout)
        self.conv2, self.bn2 = nn.Conv2d(cout, cout, 3, padding=1), nn.BatchNorm2d(cout)

    def forward(self, x):
        return F.relu(self.bn2(self.conv2(F.relu(self.bn1(self.conv1(x))))))


class TorchNet(nn.Module):
    """Five encoder blocks, four diagonal decoder blocks over cat([skip, up(below)]), 1x1 head (eval path only)."""

    def __init__(self, ncls=3):
        super().__init__()
        self.conv0_0 = Block(3, W[0])
        for i in range(1, 5):
            setattr(self, f"conv{i}_0", Block(W[i - 1], W[i]))
        for i in (3, 2, 1, 0):
            setattr(self, f"conv{i}_{4 - i}", Block(W[i] + W[i + 1], W[i]))
        self.final = nn.Conv2d(W[0], ncls, 1)

    def forward(self, x):
        up = lambda t: F.interpolate(t, scale_factor=2, mode="bilinear", align_corners=True)
        x0 = self.conv0_0(x)
        x1 = self.conv1_0(F.max_pool2d(x0, 2))
        x2 = self.conv2_0(F.max_pool2d(x1, 2))
        x3 = self.conv3_0(F.max_pool2d(x2, 2))
        x4 = self.conv4_0(F.max_pool2d(x3, 2))
        d3 = self.conv3_1(torch.cat([x3, up(x4)], 1))
        d2 = self.conv2_2(torch.cat([x2, up(d3)], 1))
        d1 = self.conv1_3(torch.cat([x1, up(d2)], 1))
        d0 = self.conv0_4(torch.cat([x0, up(d1)], 1))
        return self.final(d0)


def timed(fn, reps):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    torch.cuda.set_device(0)
    sd = S.make_state_dict(0)
    net = TorchNet().cuda().eval()
    missing = net.load_state_dict({k: v for k, v in sd.items() if not k.startswith("ds")}, strict=True)
    x = torch.rand(B, 3, 512, 512, device="cuda")
    gflop = 111.636 * B
    rows = []

    def add(name, ms, y=None, ref=None):
        r = {"impl": name, "ms_per_batch": round(ms, 3), "frames_per_s": round(B / ms * 1e3, 1), "tflops": round(gflop / ms, 1)}
        if y is not None and ref is not None:
            r["max_rel_err_vs_fp32"] = float((y.float() - ref).abs().max() / ref.abs().max())
            r["argmax_agreement"] = float((y.argmax(1) == ref.argmax(1)).float().mean())
        rows.append(r)
        print(r, file=sys.stderr)

    with torch.no_grad():
        torch.backends.cudnn.allow_tf32 = False
        torch.backends.cuda.matmul.allow_tf32 = False
        torch.backends.cudnn.benchmark = True
        ref = net(x)
        add("torch eager fp32 (TF32 off, cuDNN)", timed(lambda: net(x), max(reps // 3, 2)))
        torch.backends.cudnn.allow_tf32 = True
        y = net(x)
        add("torch eager fp32 (TF32 on)", timed(lambda: net(x), reps), y, ref)
        netc = net.to(memory_format=torch.channels_last)
        xc = x.to(memory_format=torch.channels_last)
        for dt, nm in ((torch.bfloat16, "bf16"), (torch.float16, "fp16")):
            with torch.autocast("cuda", dtype=dt):
                y = netc(xc)
                add(f"torch eager {nm} autocast, channels_last", timed(lambda: netc(xc), reps), y, ref)
            nh = TorchNet().cuda().eval()
            nh.load_state_dict(net.state_dict())
            nh = nh.to(dt).to(memory_format=torch.channels_last)
            xh = xc.to(dt)
            y = nh(xh)
            add(f"torch eager pure {nm}, channels_last", timed(lambda: nh(xh), reps), y, ref)
            del nh
        ours = NestedUNet(3).cuda().eval()
        ours.load_state_dict(sd, strict=True)
        y = ours(x)
        add("unet_b200 NestedUNet.forward (fp32 NCHW in/out, layout conversion included)", timed(lambda: ours(x), reps), y, ref)
    print(json.dumps({"batch": B, "net": "512x512", "gflop_per_batch": gflop, "rows": rows}))


if __name__ == "__main__":
    main()
