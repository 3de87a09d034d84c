"""ORACLE (test infrastructure, not product code).

numpy restatement of the OpenCV / numpy arithmetic the reference's hot path relies on.
OpenCV is NOT vendored by the reference (requirements.txt:3 `opencv-python>=4.5.0`,
unpinned); these restate the published algorithms of opencv 4.13.0 (the installed
build) and are pinned bit-exactly against cv2 itself in tests/test_oracle_cvnp.py on
seeded inputs, and against the reference's own outputs in tests/golden/.  They are the
specification the CUDA kernels in unet-_b200/csrc follow, one function per kernel.

Call sites in the reference (anchor for each function):
  bgr2gray            infer_two_stage_burr.py:317
  resize_linear       infer_two_stage_burr.py:124, :280
  resize_nearest      infer_two_stage_burr.py:307-308
  gaussian5           infer_two_stage_burr.py:85
  canny               infer_two_stage_burr.py:86, infer_enhanced_burr.py:90
  sobel_edges         infer_enhanced_burr.py:93-97
  laplacian_edges     infer_enhanced_burr.py:100-102
  ellipse / dilate / erode / morph_close / morph_open
                      infer_two_stage_burr.py:78-80,92-97, infer_enhanced_burr.py:82-84,112-116
  ccl_stats           infer_two_stage_burr.py:100, infer_enhanced_burr.py:119
  clahe / sharpen     infer_enhanced_burr.py:50-51, :61-64
"""
from __future__ import annotations

import numpy as np


# ------------------------------------------------------------------ colour / resize
def bgr2gray(bgr: np.ndarray) -> np.ndarray:
    b = bgr[..., 0].astype(np.int64)
    g = bgr[..., 1].astype(np.int64)
    r = bgr[..., 2].astype(np.int64)
    return ((3735 * b + 19235 * g + 9798 * r + (1 << 14)) >> 15).astype(np.uint8)


def linear_coeffs(src: int, dst: int, is_x: bool):
    """Per-axis source index and 11-bit weights of cv2.resize(INTER_LINEAR) on uint8.
    Returns (i0, i1, c0, c1) int32 arrays of length dst."""
    scale = np.float64(src) / np.float64(dst)
    d = np.arange(dst, dtype=np.float64)
    f = ((d + 0.5) * scale - 0.5).astype(np.float32)
    s = np.floor(f).astype(np.int32)
    f = (f - s.astype(np.float32)).astype(np.float32)
    if is_x:
        lo = s < 0
        s[lo] = 0
        f[lo] = 0
        hi = s >= src - 1
        s[hi] = src - 1
        f[hi] = 0
        i0 = s
        i1 = np.minimum(s + 1, src - 1)
    else:
        i0 = np.clip(s, 0, src - 1)
        i1 = np.clip(s + 1, 0, src - 1)
    one = np.float32(1.0)
    k = np.float32(2048.0)
    c0 = np.rint((one - f) * k).astype(np.int32)
    c1 = np.rint(f * k).astype(np.int32)
    return i0.astype(np.int32), i1.astype(np.int32), c0, c1


def resize_linear(img: np.ndarray, dsize) -> np.ndarray:
    """cv2.resize(img, (dw,dh), INTER_LINEAR) for uint8 HxW[xC]."""
    dw, dh = dsize
    sh, sw = img.shape[:2]
    x0, x1, cx0, cx1 = linear_coeffs(sw, dw, True)
    y0, y1, cy0, cy1 = linear_coeffs(sh, dh, False)
    a = img.astype(np.int32)
    if a.ndim == 2:
        a = a[..., None]
    hrow = a[:, x0, :] * cx0[None, :, None] + a[:, x1, :] * cx1[None, :, None]   # [sh,dw,C]
    top = hrow[y0] >> 4
    bot = hrow[y1] >> 4
    out = (((cy0[:, None, None] * top) >> 16) + ((cy1[:, None, None] * bot) >> 16) + 2) >> 2
    out = out.astype(np.uint8)
    return out[..., 0] if img.ndim == 2 else out


def nearest_index(src: int, dst: int) -> np.ndarray:
    # cv2 INTER_NEAREST: sx = min(floor(x * (src/dst)), src-1) with double scale
    scale = np.float64(src) / np.float64(dst)
    return np.minimum(np.floor(np.arange(dst) * scale).astype(np.int32), src - 1)


def resize_nearest(img: np.ndarray, dsize) -> np.ndarray:
    dw, dh = dsize
    ys = nearest_index(img.shape[0], dh)
    xs = nearest_index(img.shape[1], dw)
    return img[ys][:, xs]


# ------------------------------------------------------------------ borders
def _idx_reflect101(n: int, lo: int, hi: int) -> np.ndarray:
    i = np.arange(lo, n + hi)
    i = np.where(i < 0, -i, i)
    i = np.where(i >= n, 2 * (n - 1) - i, i)
    return i


def _pad(img, r, mode):
    h, w = img.shape
    if mode == "reflect101":
        yi = _idx_reflect101(h, -r, r)
        xi = _idx_reflect101(w, -r, r)
    else:  # replicate
        yi = np.clip(np.arange(-r, h + r), 0, h - 1)
        xi = np.clip(np.arange(-r, w + r), 0, w - 1)
    return img[yi][:, xi]


# ------------------------------------------------------------------ filters
GAUSS5 = np.array([14, 62, 104, 62, 14], np.int64)   # sigma 1.0, 8-bit fixed point, sum 256


def gaussian5(gray: np.ndarray) -> np.ndarray:
    """cv2.GaussianBlur(gray, (5,5), 1.0) on uint8: separable fixed-point kernel, both
    passes accumulated un-rounded, one final rounding."""
    p = _pad(gray.astype(np.int64), 2, "reflect101")
    h, w = gray.shape
    hor = sum(GAUSS5[k] * p[:, k:k + w] for k in range(5))
    acc = sum(GAUSS5[k] * hor[k:k + h, :] for k in range(5))
    return ((acc + (1 << 15)) >> 16).astype(np.uint8)


def sobel3(gray: np.ndarray, border: str):
    p = _pad(gray.astype(np.int32), 1, border)
    h, w = gray.shape
    v = lambda dy, dx: p[1 + dy:1 + dy + h, 1 + dx:1 + dx + w]
    gx = (v(-1, 1) + 2 * v(0, 1) + v(1, 1)) - (v(-1, -1) + 2 * v(0, -1) + v(1, -1))
    gy = (v(1, -1) + 2 * v(1, 0) + v(1, 1)) - (v(-1, -1) + 2 * v(-1, 0) + v(-1, 1))
    return gx, gy


def canny_candidates(img: np.ndarray, low: int, high: int) -> np.ndarray:
    """First half of cv2.Canny(img, low, high) (aperture 3, L1 gradient):
    returns u8 map 0 = not an edge, 1 = weak candidate, 2 = strong."""
    gx, gy = sobel3(img, "replicate")
    h, w = img.shape
    mag = np.abs(gx) + np.abs(gy)
    mp = np.zeros((h + 2, w + 2), np.int32)
    mp[1:-1, 1:-1] = mag
    m = lambda dy, dx: mp[1 + dy:1 + dy + h, 1 + dx:1 + dx + w]
    ax = np.abs(gx).astype(np.int64)
    ay = np.abs(gy).astype(np.int64) << 15
    tg22 = ax * 13573
    tg67 = tg22 + (ax << 16)
    horiz = ay < tg22
    vert = ay > tg67
    s_neg = (gx ^ gy) < 0
    c = m(0, 0)
    keep_h = (c > m(0, -1)) & (c >= m(0, 1))
    keep_v = (c > m(-1, 0)) & (c >= m(1, 0))
    # diagonal: s = -1 if sign differs else +1; compare mag[y-1][x-s], mag[y+1][x+s]
    d_pos = (c > m(-1, -1)) & (c > m(1, 1))
    d_neg = (c > m(-1, 1)) & (c > m(1, -1))
    keep_d = np.where(s_neg, d_neg, d_pos)
    keep = np.where(horiz, keep_h, np.where(vert, keep_v, keep_d))
    cand = (c > low) & keep
    out = np.zeros((h, w), np.uint8)
    out[cand] = 1
    out[cand & (c > high)] = 2
    return out


def _label8(fg: np.ndarray) -> np.ndarray:
    """8-connected labelling (arbitrary numbering) by a two-pass union-find on runs."""
    from scipy import ndimage
    lab, _ = ndimage.label(fg, structure=np.ones((3, 3), np.int32))
    return lab


def canny(img: np.ndarray, low: int, high: int) -> np.ndarray:
    """cv2.Canny: hysteresis = every candidate whose 8-connected component of
    candidates contains a strong pixel."""
    cand = canny_candidates(img, low, high)
    lab = _label8(cand > 0)
    strong_labels = np.unique(lab[cand == 2])
    keep = np.zeros(lab.max() + 1, bool)
    keep[strong_labels] = True
    keep[0] = False
    return np.where(keep[lab], 255, 0).astype(np.uint8)


def sobel_edges(gray: np.ndarray, thresh: int = 50) -> np.ndarray:
    """infer_enhanced_burr.py:93-97: f64 Sobel magnitude normalised by the frame's
    max, truncating uint8 cast, threshold.  A constant frame (max 0 -> NaN in the
    reference) is defined as no edges."""
    gx, gy = sobel3(gray, "reflect101")
    s2 = gx.astype(np.int64) ** 2 + gy.astype(np.int64) ** 2
    mag = np.sqrt(s2.astype(np.float64))
    mx = mag.max()
    if mx == 0:
        return np.zeros(gray.shape, np.uint8)
    m8 = (mag / mx * 255).astype(np.uint8)
    return np.where(m8 > thresh, 255, 0).astype(np.uint8)


def laplacian_edges(gray: np.ndarray, thresh: int = 15) -> np.ndarray:
    """infer_enhanced_burr.py:100-102: 4-neighbour Laplacian (REFLECT_101); the uint8
    cast of |lap| wraps modulo 256."""
    p = _pad(gray.astype(np.int32), 1, "reflect101")
    h, w = gray.shape
    v = lambda dy, dx: p[1 + dy:1 + dy + h, 1 + dx:1 + dx + w]
    lap = v(-1, 0) + v(1, 0) + v(0, -1) + v(0, 1) - 4 * v(0, 0)
    l8 = (np.abs(lap) & 255).astype(np.uint8)
    return np.where(l8 > thresh, 255, 0).astype(np.uint8)


# ------------------------------------------------------------------ morphology
def ellipse(k: int) -> np.ndarray:
    """cv2.getStructuringElement(MORPH_ELLIPSE, (k,k)) (OpenCV morph.dispatch.cpp):
    r = k//2, c = k//2, inv_r2 = 1/r^2; row i: dy = i - r; if |dy| <= r:
    dx = saturate_int(c * sqrt((r^2 - dy^2) * inv_r2)); j1 = max(c-dx,0),
    j2 = min(c+dx+1, k)."""
    r = k // 2
    c = k // 2
    inv_r2 = 1.0 / (r * r) if r else 0.0
    out = np.zeros((k, k), np.uint8)
    for i in range(k):
        dy = i - r
        if abs(dy) <= r:
            dx = int(np.rint(c * np.sqrt((r * r - dy * dy) * inv_r2)))
            j1 = max(c - dx, 0)
            j2 = min(c + dx + 1, k)
            out[i, j1:j2] = 1
    return out


def _morph(img: np.ndarray, se: np.ndarray, is_dilate: bool) -> np.ndarray:
    kh, kw = se.shape
    ay, ax = kh // 2, kw // 2
    h, w = img.shape
    ident = 0 if is_dilate else 255
    p = np.full((h + kh, w + kw), ident, np.uint8)
    p[ay:ay + h, ax:ax + w] = img
    out = np.full((h, w), ident, np.uint8)
    op = np.maximum if is_dilate else np.minimum
    for i in range(kh):
        for j in range(kw):
            if se[i, j]:
                out = op(out, p[i:i + h, j:j + w])
    return out


def dilate(img, se):
    return _morph(img, se, True)


def erode(img, se):
    return _morph(img, se, False)


def morph_close(img, se):
    return erode(dilate(img, se), se)


def morph_open(img, se):
    return dilate(erode(img, se), se)


# ------------------------------------------------------------------ CCL
def ccl_stats(img: np.ndarray):
    """cv2.connectedComponentsWithStats(img, connectivity=8): labels int32 numbered in
    order of first appearance when scanning 2x2 blocks in raster order (inside a block:
    (0,0),(0,1),(1,0),(1,1)); stats int32 [n,5] = left, top, width, height, area with
    row 0 = background bounding box / area."""
    fg = img != 0
    h, w = fg.shape
    lab = _label8(fg)
    n = int(lab.max())
    ys, xs = np.nonzero(fg)
    key = (((ys >> 1) * ((w + 1) >> 1) + (xs >> 1)) << 2) | ((ys & 1) << 1) | (xs & 1)
    l = lab[ys, xs]
    first = np.full(n + 1, np.iinfo(np.int64).max, np.int64)
    np.minimum.at(first, l, key)
    order = np.argsort(first[1:], kind="stable")
    remap = np.zeros(n + 1, np.int32)
    remap[order + 1] = np.arange(1, n + 1, dtype=np.int32)
    labels = remap[lab].astype(np.int32)
    stats = np.zeros((n + 1, 5), np.int32)
    for i in range(n + 1):
        yy, xx = np.nonzero(labels == i)
        if len(yy) == 0:
            # OpenCV leaves an empty background as left/top = INT_MAX clipped; not hit in tests
            continue
        stats[i] = (xx.min(), yy.min(), xx.max() - xx.min() + 1, yy.max() - yy.min() + 1, len(yy))
    return n + 1, labels, stats


def filter_components(labels, stats, min_area, max_area, max_ratio_int, min_side):
    """Component filter of infer_two_stage_burr.py:103-117 / infer_enhanced_burr.py:121-136
    in exact integer form: max(w,h)/(min(w,h)+1e-6) < R  <=>  max(w,h) <= R*min(w,h)
    for positive integers (the 1e-6 only breaks the tie at equality towards `<`)."""
    keep = np.zeros(len(stats), bool)
    for i in range(1, len(stats)):
        _, _, w, h, a = stats[i]
        keep[i] = (min_area <= a <= max_area and max(w, h) <= max_ratio_int * min(w, h)
                   and w >= min_side and h >= min_side)
    return keep[labels].astype(np.uint8)


# ------------------------------------------------------------------ enhance pieces
def clahe_u8(src: np.ndarray, clip_limit: float = 3.0, tiles=(8, 8)) -> np.ndarray:
    """cv2.createCLAHE(clip_limit, tiles).apply(src) for uint8 whose size divides evenly
    into the tile grid (800x448 with 8x8 tiles does)."""
    h, w = src.shape
    tx, ty = tiles
    tw, th = w // tx, h // ty
    assert tw * tx == w and th * ty == h
    area = tw * th
    clip = max(int(clip_limit * area / 256), 1)
    scale = np.float32(255.0) / np.float32(area)
    luts = np.zeros((ty, tx, 256), np.uint8)
    for j in range(ty):
        for i in range(tx):
            hist = np.bincount(src[j * th:(j + 1) * th, i * tw:(i + 1) * tw].ravel(),
                               minlength=256).astype(np.int32)
            excess = int(np.maximum(hist - clip, 0).sum())
            hist = np.minimum(hist, clip)
            hist += excess // 256
            resid = excess - (excess // 256) * 256
            if resid:
                step = max(256 // resid, 1)
                idx = np.arange(0, 256, step)[:resid]
                hist[idx] += 1
            cs = np.cumsum(hist).astype(np.float32)
            luts[j, i] = np.clip(np.rint(cs * scale), 0, 255).astype(np.uint8)
    inv_tw = np.float32(1.0) / np.float32(tw)
    inv_th = np.float32(1.0) / np.float32(th)
    xs = np.arange(w, dtype=np.float32) * inv_tw - np.float32(0.5)
    ys = np.arange(h, dtype=np.float32) * inv_th - np.float32(0.5)
    tx1 = np.floor(xs).astype(np.int32)
    ty1 = np.floor(ys).astype(np.int32)
    xa = (xs - tx1.astype(np.float32)).astype(np.float32)
    ya = (ys - ty1.astype(np.float32)).astype(np.float32)
    tx2 = np.minimum(tx1 + 1, tx - 1)
    ty2 = np.minimum(ty1 + 1, ty - 1)
    tx1 = np.maximum(tx1, 0)
    ty1 = np.maximum(ty1, 0)
    v = src.astype(np.int64)
    Y1, Y2 = ty1[:, None], ty2[:, None]
    X1, X2 = tx1[None, :], tx2[None, :]
    f = lambda Y, X: luts[Y, X, v].astype(np.float32)
    one = np.float32(1.0)
    xa2, ya2 = xa[None, :], ya[:, None]
    res = ((f(Y1, X1) * (one - xa2) + f(Y1, X2) * xa2) * (one - ya2)
           + (f(Y2, X1) * (one - xa2) + f(Y2, X2) * xa2) * ya2).astype(np.float32)
    return np.clip(np.rint(res), 0, 255).astype(np.uint8)


def sharpen3(bgr: np.ndarray) -> np.ndarray:
    """cv2.filter2D(img, -1, [[-1,-1,-1],[-1,9,-1],[-1,-1,-1]]) on uint8, REFLECT_101."""
    out = np.empty_like(bgr)
    for c in range(bgr.shape[2]):
        p = _pad(bgr[..., c].astype(np.int32), 1, "reflect101")
        h, w = bgr.shape[:2]
        acc = 10 * p[1:1 + h, 1:1 + w]
        for dy in range(3):
            for dx in range(3):
                acc = acc - p[dy:dy + h, dx:dx + w]
        out[..., c] = np.clip(acc, 0, 255).astype(np.uint8)
    return out


# ----------------------------------------------------------------------------------------------
# cv2.addWeighted on uint8 (visualize_two_stage, infer_two_stage_burr.py:141,152-154): OpenCV's SIMD
# kernel is fma(a, alpha, fma(b, beta, gamma)) in binary32, then round-half-even and saturate.
# With gamma = 0 the inner fma is the rounded product b*beta.  Pinned on all 256x256 operand pairs
# for the three weight pairs the reference uses (tests/test_oracle_cvnp.py).
# ----------------------------------------------------------------------------------------------
def add_weighted_u8(a: np.ndarray, alpha: float, b: np.ndarray, beta: float) -> np.ndarray:
    al, be = np.float32(alpha), np.float32(beta)
    inner = (b.astype(np.float64) * np.float64(be)).astype(np.float32)            # fma(b, beta, 0)
    v = (a.astype(np.float64) * np.float64(al) + inner.astype(np.float64)).astype(np.float32)   # one rounding = fma
    return np.clip(np.rint(v), 0, 255).astype(np.uint8)


OVERLAY_COLORS = ((0, 255, 0), (255, 0, 0), (255, 0, 255))     # cable, tape, burr in BGR (CLASS_COLORS[1..3], :22-26)


def overlay_blend(frame: np.ndarray, mask_cable, mask_tape, mask_burr, roi) -> np.ndarray:
    """The four addWeighted steps of visualize_two_stage (infer_two_stage_burr.py:134-154; the same lines in
    infer_enhanced_burr.py:157-175): dim outside the ROI, then tint cable / tape / burr."""
    x1, y1, x2, y2 = roi
    dimmed = frame.copy()
    outside = np.ones(frame.shape[:2], bool)
    outside[y1:y2, x1:x2] = False
    dimmed[outside] = 0
    out = add_weighted_u8(frame, 0.7, dimmed, 0.3)
    for mask, colour, a, b in ((mask_cable, OVERLAY_COLORS[0], 0.6, 0.4), (mask_tape, OVERLAY_COLORS[1], 0.6, 0.4),
                               (mask_burr, OVERLAY_COLORS[2], 0.5, 0.5)):
        tint = np.zeros_like(out)
        tint[mask > 0] = colour
        out = add_weighted_u8(out, a, tint, b)
    return out


# ----------------------------------------------------------------------------------------------
# cv2.GaussianBlur(v.reshape(-1,1), (1,31), sigmaX=0) on float32 (src/utils/geometry_enhanced.py:70-76): kernel =
# float32(getGaussianKernel(31, sigma=5)), BORDER_REFLECT_101, and OpenCV's symmetric column filter order
#   s = k0*c;  s += kj * (p[+j] + p[-j])  for j = 1..15,   every product and sum rounded to binary32 (no FMA).
# ----------------------------------------------------------------------------------------------
def gauss31_kernel_f32() -> np.ndarray:
    x = np.arange(31, dtype=np.float64) - 15.0
    t = np.exp(-0.5 / 25.0 * x * x)
    return (t / t.sum()).astype(np.float32)


def gauss31_f32(v: np.ndarray) -> np.ndarray:
    v = np.asarray(v, np.float32)
    n = len(v)
    k = gauss31_kernel_f32()
    idx = np.abs(np.arange(-15, n + 15))
    idx = np.where(idx >= n, 2 * (n - 1) - idx, idx)
    p = v[idx]
    s = (k[15] * p[15:15 + n]).astype(np.float32)
    for j in range(1, 16):
        pair = (p[15 + j:15 + j + n] + p[15 - j:15 - j + n]).astype(np.float32)
        s = (s + (k[15 + j] * pair).astype(np.float32)).astype(np.float32)
    return s
