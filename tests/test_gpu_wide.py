"""GPU parity of the stage-2 kernels on WIDE and LARGE frames: several strips per row in the column-walking kernels (a warp
covers 30 words: 120 pixels in k_blur5_cols, 960 in k_morph_cols / k_burr_chain_cols), more than eight strips per row (a second
block column), frames whose bit planes do not fit the per-frame hysteresis flood (the union-find path), 1080p source frames.
Everything bit-exact against cv2 / the oracle, through the C ABI."""
import cv2
import numpy as np
import pytest
import torch

import unet_b200.synth as S
from oracle import ref_pipeline as R

pytestmark = pytest.mark.gpu


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def same(name, got, want):
    got = got.cpu().numpy() if isinstance(got, torch.Tensor) else got
    assert got.shape == want.shape, (name, got.shape, want.shape)
    bad = np.argwhere(got != want)
    assert len(bad) == 0, f"{name}: {len(bad)} of {want.size} differ; first at {bad[0].tolist()}"


def gray_frames(n, h, w, seed):
    fr = np.concatenate([S.frames_cable(n - 1, h, w, seed=seed), S.frames_noise(1, h, w, seed=seed + 1)])
    return np.stack([cv2.cvtColor(f, cv2.COLOR_BGR2GRAY) for f in fr])


@pytest.mark.parametrize("hw", [(40, 1100), (300, 1920), (1080, 1920), (64, 2048), (23, 964)])
def test_blur_canny_wide(hw):
    """(40,1100): ten blur strips = two block columns; (300,1920): flood hysteresis at 60 words per row; (1080,1920): the bit
    planes exceed the flood's shared memory -> union-find hysteresis; (23,964): a last strip of one word."""
    from unet_b200 import _lib
    h, w = hw
    gray = gray_frames(2, h, w, seed=h + w)
    c = _lib.Context(0, max_batch=2, frame_hw=hw, net_hw=(16, 16))
    try:
        blur = np.stack([cv2.GaussianBlur(g, (5, 5), 1.0) for g in gray])
        same("gauss", c.gaussian5(dev(gray)), blur)
        want = np.stack([cv2.Canny(b, 50, 150) for b in blur])
        same("blur_canny", c.blur_canny(dev(gray), 50, 150), want)
        same("canny", c.canny(dev(blur), 50, 150), want)
    finally:
        c.close()


@pytest.mark.parametrize("hw", [(64, 1100), (200, 1920), (70, 2000)])
@pytest.mark.parametrize("k", [3, 8, 25, 7])
def test_morphology_wide(hw, k):
    """Two and three strips per row; k = 7 takes the run-time footprint (no compile-time table)."""
    from unet_b200 import _lib
    h, w = hw
    se = cv2.getStructuringElement(cv2.MORPH_ELLIPSE, (k, k))
    b = np.stack([S.mask_blobs(h, w, seed=k + i, fill=f) for i, f in enumerate((0.3, 0.02))])
    c = _lib.Context(0, max_batch=2, frame_hw=hw, net_hw=(16, 16))
    try:
        d = dev(b)
        same(f"dilate{k}", c.morph("dilate", k, d), np.stack([cv2.dilate(x, se) for x in b]))
        same(f"erode{k}", c.morph("erode", k, d), np.stack([cv2.erode(x, se) for x in b]))
        same(f"close{k}", c.morph("close", k, d), np.stack([cv2.morphologyEx(x, cv2.MORPH_CLOSE, se) for x in b]))
        same(f"open{k}", c.morph("open", k, d), np.stack([cv2.morphologyEx(x, cv2.MORPH_OPEN, se) for x in b]))
    finally:
        c.close()


@pytest.mark.parametrize("hw", [(120, 1100), (448, 1920)])
def test_burr_stage_wide(hw):
    """The whole burr stage (both variants; the chain kernel with two strips per row) against the oracle, plus a parameter set
    without compile-time tables (band 7, close 3, open 3)."""
    from unet_b200 import _lib
    h, w = hw
    gray = gray_frames(2, h, w, seed=7)
    masks = np.stack([S.mask_cable_synth(h, w), S.mask_blobs(h, w, 3, 0.01)])
    c = _lib.Context(0, max_batch=2, frame_hw=hw, net_hw=(16, 16))
    try:
        out, px = c.burr(dev(gray), dev(masks), 0)
        oute, _ = c.burr(dev(gray), dev(masks), 1)
        for i in range(2):
            same(f"std[{i}]", out[i], R.burrs_standard(gray[i], masks[i]))
            same(f"enh[{i}]", oute[i], R.burrs_enhanced(gray[i], masks[i]))
        assert px.cpu().tolist() == [int(out[i].sum()) for i in range(2)]
        p = _lib.standard_params()
        p.band_k, p.close_k, p.open_k = 7, 3, 3
        out7, _ = c.burr(dev(gray), dev(masks), 0, p)
        for i in range(2):
            E = lambda k: cv2.getStructuringElement(cv2.MORPH_ELLIPSE, (k, k))
            edges = cv2.Canny(cv2.GaussianBlur(gray[i], (5, 5), 1.0), p.canny_low, p.canny_high)
            band = cv2.dilate(masks[i], E(7)) & (1 - masks[i])
            cand = cv2.morphologyEx(cv2.morphologyEx(((edges > 0) & (band > 0)).astype(np.uint8), cv2.MORPH_CLOSE, E(3)), cv2.MORPH_OPEN, E(3))
            n, lab, st, _ = cv2.connectedComponentsWithStats(cand, connectivity=8)
            want = np.zeros_like(cand)
            for j in range(1, n):
                x, y, cw, ch, area = st[j]
                if p.min_area <= area <= p.max_area and max(cw, ch) <= p.max_aspect * min(cw, ch) and cw >= p.min_side and ch >= p.min_side:
                    want[lab == j] = 1
            same(f"band7[{i}]", out7[i], want)
    finally:
        c.close()


def test_properties_at_full_batch():
    """Size-independent properties on a full batch of 64 frames of 800x448 (the bench's shape): CLOSE and OPEN with odd ellipses are idempotent,
    dilation and erosion are dual under complement (inside the image: out-of-image samples are ignored by both), the blur of a
    constant frame is that constant, hysteresis output lies inside the candidates of a lower threshold pair."""
    from unet_b200 import _lib
    B, h, w = 64, 448, 800
    c = _lib.Context(0, max_batch=B, frame_hw=(h, w), net_hw=(16, 16))
    try:
        m = dev(np.stack([S.mask_blobs(h, w, seed=i, fill=0.05 + 0.005 * i) for i in range(B)]))
        for k in (3, 5, 25):                   # odd sizes: with cv2's even ellipses (anchor off centre) CLOSE is not idempotent either
            cl = c.morph("close", k, m)
            assert torch.equal(c.morph("close", k, cl), cl), k
            op = c.morph("open", k, m)
            assert torch.equal(c.morph("open", k, op), op), k
            assert bool((cl >= m).all()) and bool((op <= m).all())
        for k in (3, 5):                       # odd sizes: symmetric footprints, so erode(m) == 1 - dilate(1 - m) up to the border rule
            er = c.morph("erode", k, m)
            du = 1 - c.morph("dilate", k, 1 - m)
            r = k // 2
            assert torch.equal(er[:, r:-r, r:-r], du[:, r:-r, r:-r]), k
        const = torch.full((B, h, w), 77, dtype=torch.uint8, device="cuda")
        assert torch.equal(c.gaussian5(const), const)
        g = dev(gray_frames(4, h, w, seed=3)).repeat(16, 1, 1)
        e_hi = c.blur_canny(g, 50, 150)
        e_lo = c.blur_canny(g, 30, 100)
        assert bool(((e_hi > 0) <= (e_lo > 0)).all())          # raising both thresholds can only remove edge pixels
        assert torch.equal(e_hi[:4], e_hi[4:8])                   # identical frames give identical maps at every batch position
    finally:
        c.close()


@pytest.mark.parametrize("hw", [(37, 53), (5, 7), (448, 800), (33, 65)])
def test_gray_odd_sizes_and_unaligned_frames(hw):
    """BGR2GRAY (16 pixels per thread with 16-byte vectors when aligned): pixel counts that are not multiples of 16, and a frame
    pointer one byte off alignment (the scalar path), against cv2."""
    from unet_b200 import _lib
    h, w = hw
    fr = S.frames_noise(3, h, w, seed=h * w)
    c = _lib.Context(0, max_batch=3, frame_hw=hw, net_hw=(16, 16))
    try:
        _, g = c.preprocess(dev(fr), want_x=False, want_gray=True)
        for i in range(3):
            same(f"gray[{i}]", g[i], cv2.cvtColor(fr[i], cv2.COLOR_BGR2GRAY))
        raw = torch.zeros(fr.size + 1, dtype=torch.uint8, device="cuda")
        raw[1:] = dev(fr).reshape(-1)
        _, g = c.preprocess(raw[1:].view(3, h, w, 3), want_x=False, want_gray=True)
        for i in range(3):
            same(f"gray unaligned[{i}]", g[i], cv2.cvtColor(fr[i], cv2.COLOR_BGR2GRAY))
    finally:
        c.close()
