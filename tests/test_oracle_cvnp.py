"""Pins oracle/cvnp.py (numpy restatement of OpenCV's arithmetic = the kernel spec)
bit-exactly against the installed cv2 on seeded inputs, including ragged sizes."""
import cv2
import numpy as np
import pytest

import unet_b200.synth as S
from oracle import cvnp

SIZES = [(448, 800), (37, 53), (64, 96)]


def _frame(kind, h, w, seed):
    return (S.frames_cable if kind == "cable" else S.frames_noise)(1, h, w, seed=seed)[0]


@pytest.mark.parametrize("kind", ["cable", "noise"])
@pytest.mark.parametrize("hw", SIZES)
def test_filters(kind, hw):
    h, w = hw
    fr = _frame(kind, h, w, 11)
    gray = cv2.cvtColor(fr, cv2.COLOR_BGR2GRAY)
    assert np.array_equal(cvnp.bgr2gray(fr), gray)
    blur = cv2.GaussianBlur(gray, (5, 5), 1.0)
    assert np.array_equal(cvnp.gaussian5(gray), blur)
    for lo, hi in ((50, 150), (30, 100)):
        assert np.array_equal(cvnp.canny(blur, lo, hi), cv2.Canny(blur, lo, hi))
    sx = cv2.Sobel(gray, cv2.CV_64F, 1, 0, ksize=3)
    sy = cv2.Sobel(gray, cv2.CV_64F, 0, 1, ksize=3)
    mag = np.sqrt(sx ** 2 + sy ** 2)
    m8 = np.uint8(mag / mag.max() * 255)
    assert np.array_equal(cvnp.sobel_edges(gray), cv2.threshold(m8, 50, 255, cv2.THRESH_BINARY)[1])
    lap = np.abs(cv2.Laplacian(gray, cv2.CV_64F)).astype(np.uint8)
    assert np.array_equal(cvnp.laplacian_edges(gray), cv2.threshold(lap, 15, 255, cv2.THRESH_BINARY)[1])


@pytest.mark.parametrize("src,dst", [((448, 800), (512, 512)), ((1080, 1920), (448, 800)),
                                     ((1920, 1080), (448, 800)), ((100, 37), (64, 64))])
def test_resize(src, dst):
    fr = S.frames_noise(1, src[0], src[1], seed=3)[0]
    want = cv2.resize(fr, (dst[1], dst[0]), interpolation=cv2.INTER_LINEAR)
    assert np.array_equal(cvnp.resize_linear(fr, (dst[1], dst[0])), want)
    m = S.mask_blobs(dst[0], dst[1], seed=1, fill=0.5)
    want = cv2.resize(m, (src[1], src[0]), interpolation=cv2.INTER_NEAREST)
    assert np.array_equal(cvnp.resize_nearest(m, (src[1], src[0])), want)


@pytest.mark.parametrize("k", [2, 3, 5, 8, 25])
def test_morphology(k):
    se = cv2.getStructuringElement(cv2.MORPH_ELLIPSE, (k, k))
    assert np.array_equal(cvnp.ellipse(k), se)
    for h, w, fill in ((448, 800, 0.3), (50, 70, 0.05), (31, 33, 0.6)):
        b = S.mask_blobs(h, w, seed=k, fill=fill)
        assert np.array_equal(cvnp.dilate(b, se), cv2.dilate(b, se))
        assert np.array_equal(cvnp.erode(b, se), cv2.erode(b, se))
        assert np.array_equal(cvnp.morph_close(b, se), cv2.morphologyEx(b, cv2.MORPH_CLOSE, se))
        assert np.array_equal(cvnp.morph_open(b, se), cv2.morphologyEx(b, cv2.MORPH_OPEN, se))


@pytest.mark.parametrize("hw,fill", [((448, 800), 0.3), ((447, 799), 0.45), ((33, 65), 0.6),
                                     ((16, 16), 0.0), ((16, 16), 1.0)])
def test_ccl(hw, fill):
    b = S.mask_blobs(hw[0], hw[1], seed=9, fill=fill)
    n, lab, st, _ = cv2.connectedComponentsWithStats(b, connectivity=8)
    n2, lab2, st2 = cvnp.ccl_stats(b)
    assert n == n2
    assert np.array_equal(lab, lab2)
    if 0.0 < fill < 1.0:
        assert np.array_equal(st, st2)
    else:
        assert np.array_equal(st[1:], st2[1:])


def test_clahe_sharpen():
    fr = S.frames_cable(1, 448, 800, seed=2)[0]
    l = cv2.cvtColor(fr, cv2.COLOR_BGR2LAB)[..., 0]
    assert np.array_equal(cvnp.clahe_u8(l), cv2.createCLAHE(clipLimit=3.0, tileGridSize=(8, 8)).apply(l))
    k = np.array([[-1, -1, -1], [-1, 9, -1], [-1, -1, -1]])
    assert np.array_equal(cvnp.sharpen3(fr), cv2.filter2D(fr, -1, k))


def test_filter_integer_form():
    rng = np.random.default_rng(0)
    for _ in range(2000):
        w, h = int(rng.integers(1, 60)), int(rng.integers(1, 60))
        for r in (5, 6):
            assert (max(w, h) / (min(w, h) + 1e-6) < float(r)) == (max(w, h) <= r * min(w, h))


@pytest.mark.parametrize("alpha,beta", [(0.7, 0.3), (0.6, 0.4), (0.5, 0.5)])
def test_add_weighted_all_operand_pairs(alpha, beta):
    """cv2.addWeighted on u8 for the reference's three weight pairs, all 256 x 256 operand pairs, and a frame-sized
    3-channel array (OpenCV's SIMD body)."""
    a = np.repeat(np.arange(256, dtype=np.uint8)[:, None], 256, 1)
    b = np.ascontiguousarray(a.T)
    assert np.array_equal(cvnp.add_weighted_u8(a, alpha, b, beta), cv2.addWeighted(a, alpha, b, beta, 0))
    rng = np.random.default_rng(0)
    x = rng.integers(0, 256, (120, 203, 3), dtype=np.uint8)
    y = rng.integers(0, 256, (120, 203, 3), dtype=np.uint8)
    assert np.array_equal(cvnp.add_weighted_u8(x, alpha, y, beta), cv2.addWeighted(x, alpha, y, beta, 0))


@pytest.mark.parametrize("n", [448, 512, 96, 33, 1000])
def test_gauss31_float_profile_filter(n):
    """cv2.GaussianBlur(v.reshape(-1,1), (1,31), 0) on float32 (geometry_enhanced.py:70-76): kernel and operation order."""
    assert np.array_equal(cvnp.gauss31_kernel_f32(), cv2.getGaussianKernel(31, 0, cv2.CV_32F).reshape(-1))
    rng = np.random.default_rng(n)
    for trial in range(4):
        v = rng.integers(0, 800, n).astype(np.float32)
        if trial % 2:
            v[rng.random(n) < 0.4] = 0
        assert np.array_equal(cvnp.gauss31_f32(v), cv2.GaussianBlur(v.reshape(-1, 1), (1, 31), sigmaX=0).reshape(-1))


def test_gauss31_table_in_the_kernel_source():
    """The 16 hex-float constants of c_gauss31 (csrc/stage2.cu) are float32(cv2.getGaussianKernel(31, 0)), centre first."""
    import os, re
    src = open(os.path.join(os.path.dirname(__file__), "..", "unet-_b200", "csrc", "stage2.cu")).read()
    body = re.search(r"c_gauss31\[16\] = \{(.*?)\};", src, re.S).group(1)
    vals = np.array([float.fromhex(t.strip().rstrip("f")) for t in body.split(",")], np.float64)
    k = cv2.getGaussianKernel(31, 0, cv2.CV_32F).reshape(-1)
    assert np.array_equal(vals.astype(np.float32), k[15:]) and np.array_equal(vals, k[15:].astype(np.float64))
