"""Drop-in for the reference's ``NestedUNet`` module contract
(/root/reference/src/models/unetpp.py:29-135).

Same constructor signature, same 134-key ``state_dict`` (so
``load_state_dict(ckpt['model'], strict=True)`` from infer_two_stage_burr.py:214-216 works
unchanged), same eval-mode ``forward(x[B,3,H,W]) -> logits[B,num_classes,H,W]`` in fp32.
The eval forward runs entirely in libunetb200 (bf16 tcgen05 implicit-GEMM convs with BN
folded, fp32 accumulation, fp32 head); there is no PyTorch/cuDNN fallback.

Out of scope (raise NotImplementedError): training-mode forward (deep-supervision list,
unetpp.py:121-133) and ``pretrained_encoder=True`` (ResNet50 branch, unetpp.py:52-65),
neither of which the hot-path CLIs use (infer_two_stage_burr.py:214).
"""
from __future__ import annotations

import torch
import torch.nn as nn

from . import _lib

_WIDTHS = (32, 64, 128, 256, 512)


def _double_conv(cin: int, cout: int) -> nn.Module:
    """Parameter container with the reference ConvBlock's key names (unetpp.py:13-21)."""
    m = nn.Module()
    m.conv1 = nn.Conv2d(cin, cout, kernel_size=3, padding=1)
    m.bn1 = nn.BatchNorm2d(cout)
    m.conv2 = nn.Conv2d(cout, cout, kernel_size=3, padding=1)
    m.bn2 = nn.BatchNorm2d(cout)
    return m


class NestedUNet(nn.Module):
    def __init__(self, num_classes: int, input_channels: int = 3, deep_supervision: bool = True,
                 pretrained_encoder: bool = False) -> None:
        super().__init__()
        if pretrained_encoder:
            raise NotImplementedError("pretrained_encoder=True (ResNet50 branch) is outside the B200 hot path")
        if input_channels != 3:
            raise NotImplementedError("the B200 path is built for 3-channel input")
        self.deep_supervision = deep_supervision
        self.num_classes = num_classes
        w = _WIDTHS
        chans = [("conv0_0", input_channels, w[0])] + [(f"conv{i}_0", w[i - 1], w[i]) for i in range(1, 5)]
        chans += [(f"conv{i}_{4 - i}", w[i] + w[i + 1], w[i]) for i in (3, 2, 1, 0)]
        for name, cin, cout in chans:
            setattr(self, name, _double_conv(cin, cout))
        self.final = nn.Conv2d(w[0], num_classes, kernel_size=1)
        if deep_supervision:                       # present in checkpoints, never evaluated in eval mode
            self.ds3_1 = nn.Conv2d(w[3], num_classes, kernel_size=1)
            self.ds2_2 = nn.Conv2d(w[2], num_classes, kernel_size=1)
            self.ds1_3 = nn.Conv2d(w[1], num_classes, kernel_size=1)
        self._ctx = None
        self._weights_dirty = True
        self.register_load_state_dict_post_hook(lambda module, incompatible: module._invalidate())

    # ---- weight-prep invalidation (BN fold + bf16 repack happen lazily at first forward)
    def _invalidate(self):
        self._weights_dirty = True

    def _apply(self, fn, *a, **k):
        self._weights_dirty = True
        return super()._apply(fn, *a, **k)

    def train(self, mode: bool = True):
        self._weights_dirty = True
        return super().train(mode)

    def _context(self, x: torch.Tensor) -> "_lib.Context":
        b, _, h, w = x.shape
        c = self._ctx
        if (c is None or c.device != x.device or c.max_batch < b or c.net_hw[0] * c.net_hw[1] < h * w
                or c.net_hw[1] < w or c.net_hw[0] < h):
            if c is not None:
                c.close()
            self._ctx = _lib.Context(x.device.index or 0, max_batch=b, frame_hw=(16, 16), net_hw=(h, w),
                                     num_classes=self.num_classes)
            self._weights_dirty = True
        if self._weights_dirty:
            self._ctx.set_weights(self.state_dict())
            self._weights_dirty = False
            self._check_saturation = True        # the first forward with new weights counts clamped activations
        return self._ctx

    def forward(self, x: torch.Tensor):
        if self.training:
            raise NotImplementedError("training-mode forward is outside the B200 inference hot path; call .eval()")
        if not x.is_cuda:
            raise _lib.UbError("NestedUNet (unet_b200) runs on a CUDA sm_100 device only; there is no CPU fallback")
        if x.dim() != 4 or x.shape[1] != 3 or x.shape[2] % 16 or x.shape[3] % 16:
            raise RuntimeError(f"expected input [B,3,H,W] with H,W multiples of 16, got {tuple(x.shape)}")
        x = x.contiguous().float()
        with torch.cuda.device(x.device):
            ctx = self._context(x)
            y = ctx.forward(x)
            if getattr(self, "_check_saturation", False):
                self._check_saturation = False
                n = ctx.get_stat("saturated", reset=True)
                if n:
                    import warnings
                    warnings.warn(f"unet_b200.NestedUNet: {n} activations exceeded the fp16 operand range (65504) and were "
                                  f"clamped in the first forward with these weights; logits may differ from the fp32 model",
                                  RuntimeWarning, stacklevel=2)
            return y
