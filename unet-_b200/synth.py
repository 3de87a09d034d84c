"""Seeded synthetic inputs for the inspection hot path.

The reference ships no video, no checkpoint and no test data
(/root/reference/.gitignore:30-37), so every test and benchmark input is generated
here.  Pure numpy / torch-CPU; no device work.
"""
from __future__ import annotations

import numpy as np


def frames_noise(b: int, h: int = 448, w: int = 800, seed: int = 0) -> np.ndarray:
    """iid uniform uint8 BGR frames [b,h,w,3]: worst case for Canny / CCL."""
    rng = np.random.default_rng(seed)
    return rng.integers(0, 256, size=(b, h, w, 3), dtype=np.uint8)


def frames_cable(b: int, h: int = 448, w: int = 800, seed: int = 0) -> np.ndarray:
    """Structured frames: dark noisy background, a bright vertical "cable" band at
    x in [0.40w, 0.55w) with a sinusoidal edge wobble, and 20 bright blobs ("burrs")
    touching the band's outer edges."""
    rng = np.random.default_rng(seed)
    out = np.empty((b, h, w, 3), np.uint8)
    ys = np.arange(h, dtype=np.float64)[:, None]
    xs = np.arange(w, dtype=np.float64)[None, :]
    for i in range(b):
        img = rng.normal(30.0, 10.0, size=(h, w))
        phase = rng.uniform(0, 2 * np.pi)
        wob = 6.0 * np.sin(2 * np.pi * ys / 64.0 + phase)
        left = 0.40 * w + wob
        right = 0.55 * w + wob
        band = (xs >= left) & (xs < right)
        img = img + 150.0 * band
        for _ in range(20):
            sz = int(rng.integers(6, 13))
            y0 = int(rng.integers(0, max(1, h - sz)))
            side = int(rng.integers(0, 2))
            edge = (right if side else left)[y0, 0]
            x0 = int(edge) if side else int(edge) - sz
            x0 = min(max(x0, 0), w - sz)
            img[y0:y0 + sz, x0:x0 + sz] += 80.0
        img = img + rng.normal(0.0, 8.0, size=(h, w))
        for c, gain in enumerate((1.0, 0.95, 0.9)):
            out[i, :, :, c] = np.clip(img * gain, 0, 255).astype(np.uint8)
    return out


def mask_cable_synth(h: int = 448, w: int = 800, roi=None) -> np.ndarray:
    """Cable mask decoupled from the net: 1 inside the un-wobbled band (and ROI)."""
    m = np.zeros((h, w), np.uint8)
    m[:, int(0.40 * w):int(0.55 * w)] = 1
    if roi is not None:
        x1, y1, x2, y2 = roi
        keep = np.zeros_like(m)
        keep[y1:y2, x1:x2] = 1
        m &= keep
    return m


def mask_blobs(h: int, w: int, seed: int = 0, fill: float = 0.3) -> np.ndarray:
    """Random binary image {0,1}; dense, many small components."""
    rng = np.random.default_rng(seed)
    return (rng.random((h, w)) < fill).astype(np.uint8)


def pred_classes(h: int = 512, w: int = 512, seed: int = 0, tape: bool = True, defects: bool = False) -> np.ndarray:
    """Synthetic class-id map (0 background, 1 cable, 2 tape) for the geometry measurements: a wobbling cable
    band with ragged edges, a tape wrap that covers it over a range of rows and hugs it on both sides elsewhere,
    plus stray blobs and single-pixel specks of both classes."""
    rng = np.random.default_rng(seed)
    p = np.zeros((h, w), np.uint8)
    phase = rng.uniform(0, 2 * np.pi)
    y = np.arange(h)
    cx = 0.5 * w + 0.03 * w * np.sin(2 * np.pi * y / max(h / 3.0, 8.0) + phase)
    half = 0.08 * w + rng.integers(-2, 3, h)
    xs = np.arange(w)[None, :]
    cable = np.abs(xs - cx[:, None]) <= half[:, None]
    cable[: h // 12] = False
    p[cable] = 1
    if tape:
        t_half = half + 0.04 * w + rng.integers(-3, 4, h)
        wrap = np.abs(xs - cx[:, None]) <= t_half[:, None]
        rows = (y > h // 4) & (y < (3 * h) // 4)
        p[wrap & ~cable & rows[:, None]] = 2                      # tape beside the cable
        full = (y > (2 * h) // 5) & (y < h // 2)
        p[wrap & full[:, None]] = 2                               # tape over the cable: the cable splits in two
    for cls, n in ((1, 6), (2, 6)):                               # stray blobs (some above the 50 px area gate)
        for _ in range(n):
            by, bx = rng.integers(0, h - 12), rng.integers(0, w - 12)
            bh, bw = rng.integers(2, 12, 2)
            p[by:by + bh, bx:bx + bw] = cls
    specks = rng.random((h, w))
    p[specks < 0.002] = 1
    p[specks > 0.998] = 2
    if defects:                                                   # defect classes 3..6 and pin-holes in the tape
        for cls in (3, 4, 5, 6):
            for _ in range(3):
                by, bx = rng.integers(0, h - 10), rng.integers(0, w - 10)
                bh, bw = rng.integers(1, 10, 2)
                p[by:by + bh, bx:bx + bw] = cls
        p[(p == 2) & (rng.random((h, w)) < 0.03)] = 0
    return p


def make_state_dict(seed: int = 0, num_classes: int = 3, input_channels: int = 3,
                    head_gain: float = 1.0):
    """Random-init checkpoint "W0" with the reference's 134-key layout
    (src/models/unetpp.py:40-91): default-init-like conv weights, randomised BN
    statistics so BN folding is exercised.  ``head_gain`` scales final.weight ("W1")."""
    import torch
    g = torch.Generator().manual_seed(seed)
    nb = [32, 64, 128, 256, 512]
    blocks = [("conv0_0", input_channels, nb[0]), ("conv1_0", nb[0], nb[1]),
              ("conv2_0", nb[1], nb[2]), ("conv3_0", nb[2], nb[3]),
              ("conv4_0", nb[3], nb[4]), ("conv3_1", nb[3] + nb[4], nb[3]),
              ("conv2_2", nb[2] + nb[3], nb[2]), ("conv1_3", nb[1] + nb[2], nb[1]),
              ("conv0_4", nb[0] + nb[1], nb[0])]
    sd = {}

    def conv(name, cin, cout, k):
        # ReLU-preserving scale (He-uniform) so activations neither vanish nor explode.
        bound = (6.0 / (cin * k * k)) ** 0.5
        sd[name + ".weight"] = (torch.rand(cout, cin, k, k, generator=g) * 2 - 1) * bound
        sd[name + ".bias"] = (torch.rand(cout, generator=g) * 2 - 1) * 0.1

    def bn(name, c):
        sd[name + ".weight"] = torch.rand(c, generator=g) + 0.5
        sd[name + ".bias"] = torch.randn(c, generator=g) * 0.1
        sd[name + ".running_mean"] = torch.randn(c, generator=g) * 0.1
        sd[name + ".running_var"] = torch.rand(c, generator=g) + 0.5
        sd[name + ".num_batches_tracked"] = torch.tensor(1, dtype=torch.int64)

    for blk, cin, cout in blocks:
        conv(f"{blk}.conv1", cin, cout, 3)
        bn(f"{blk}.bn1", cout)
        conv(f"{blk}.conv2", cout, cout, 3)
        bn(f"{blk}.bn2", cout)
    conv("final", nb[0], num_classes, 1)
    sd["final.weight"] = sd["final.weight"] * head_gain
    conv("ds3_1", nb[3], num_classes, 1)
    conv("ds2_2", nb[2], num_classes, 1)
    conv("ds1_3", nb[1], num_classes, 1)
    return sd


def _forward_calibrating(sd, x):
    """fp32 forward of the NestedUNet topology (src/models/unetpp.py:93-119) on x that REPLACES every BatchNorm's running
    statistics in `sd` by the statistics of its own input, layer by layer (what training would have left there).
    Test-data helper: plain torch.nn.functional, not part of the product path."""
    import torch
    import torch.nn.functional as F

    def block(name, t):
        for i in (1, 2):
            t = F.conv2d(t, sd[f"{name}.conv{i}.weight"], sd[f"{name}.conv{i}.bias"], padding=1)
            mu, var = t.mean((0, 2, 3)), t.var((0, 2, 3), unbiased=False)
            sd[f"{name}.bn{i}.running_mean"], sd[f"{name}.bn{i}.running_var"] = mu.clone(), var.clone()
            t = F.relu(F.batch_norm(t, mu, var, sd[f"{name}.bn{i}.weight"], sd[f"{name}.bn{i}.bias"], False, 0.0, 1e-5))
        return t

    up = lambda t: F.interpolate(t, scale_factor=2, mode="bilinear", align_corners=True)
    pool = lambda t: F.max_pool2d(t, 2)
    x0 = block("conv0_0", x); x1 = block("conv1_0", pool(x0)); x2 = block("conv2_0", pool(x1))
    x3 = block("conv3_0", pool(x2)); x4 = block("conv4_0", pool(x3))
    x31 = block("conv3_1", torch.cat([x3, up(x4)], 1)); x22 = block("conv2_2", torch.cat([x2, up(x31)], 1))
    x13 = block("conv1_3", torch.cat([x1, up(x22)], 1)); x04 = block("conv0_4", torch.cat([x0, up(x13)], 1))
    return F.conv2d(x04, sd["final.weight"], sd["final.bias"])


def make_trained_like_state_dict(seed: int = 0, num_classes: int = 3, conv_scale: float = 3.0, gamma_max: float = 4.0,
                                 head_gain: float = 8.0):
    """A checkpoint shaped like a trained one rather than a fresh one: conv weights `conv_scale` times the
    default init, BatchNorm gammas in [0.5, gamma_max], running statistics calibrated on structured frames so every BN
    actually normalises its input (activations of order gamma, not of order 1), and a head with separated logits."""
    import torch
    sd = make_state_dict(seed, num_classes, head_gain=head_gain)
    g = torch.Generator().manual_seed(seed + 77)
    for k in list(sd):
        if ".conv" in k and k.endswith(".weight"):
            sd[k] = sd[k] * conv_scale
        if ".bn" in k and k.endswith(".weight"):
            sd[k] = 0.5 + torch.rand(sd[k].shape, generator=g) * (gamma_max - 0.5)
    fr = np.concatenate([frames_cable(2, 128, 160, seed=seed + 5), frames_speckle(1, 128, 160, seed=seed + 6, n_blobs=40)])
    x = torch.from_numpy(fr[..., ::-1].copy()).permute(0, 3, 1, 2).float() / 255.0
    with torch.no_grad():
        _forward_calibrating(sd, x)
    return sd


def make_overflow_state_dict(seed: int = 0):
    """W0 with the two BatchNorm gammas of conv0_0 scaled (x200, x1000) so that the block's output exceeds the fp16 range
    (65504) while every folded weight stays representable: the conv epilogues clamp -- what the saturation counter
    exists to report."""
    sd = make_state_dict(seed)
    sd["conv0_0.bn1.weight"] = sd["conv0_0.bn1.weight"] * 200.0
    sd["conv0_0.bn2.weight"] = sd["conv0_0.bn2.weight"] * 1000.0
    return sd


def frames_speckle(b: int, h: int = 448, w: int = 800, seed: int = 0, n_blobs: int = 400) -> np.ndarray:
    """Smooth dark background with sparse bright rectangles (3..14 px sides): yields
    many mid-sized edge components, so the component filter keeps a non-trivial subset."""
    rng = np.random.default_rng(seed)
    out = np.empty((b, h, w, 3), np.uint8)
    for i in range(b):
        img = np.full((h, w), 40.0) + rng.normal(0, 1.5, size=(h, w))
        for _ in range(n_blobs):
            bh = int(rng.integers(3, 15))
            bw = int(rng.integers(3, 15))
            y0 = int(rng.integers(0, h - bh))
            x0 = int(rng.integers(0, w - bw))
            img[y0:y0 + bh, x0:x0 + bw] += float(rng.integers(40, 200))
        for c, gain in enumerate((1.0, 0.97, 0.93)):
            out[i, :, :, c] = np.clip(img * gain, 0, 255).astype(np.uint8)
    return out


def status_frames(n_each: int = 16, h: int = 448, w: int = 800, seed: int = 100) -> np.ndarray:
    """4 * n_each frames of four kinds -- cable with burr blobs, sparse speckle, iid noise, flat frames without any
    edge -- for counting how often two implementations print the same [BURR!] / [OK] status."""
    rng = np.random.default_rng(seed)
    flat = np.clip(rng.normal(60.0, 1.0, size=(n_each, h, w, 3)), 0, 255).astype(np.uint8)
    return np.concatenate([frames_cable(n_each, h, w, seed=seed), frames_speckle(n_each, h, w, seed=seed + 1),
                           frames_noise(n_each, h, w, seed=seed + 2), flat])
