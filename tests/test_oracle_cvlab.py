"""CPU: pins oracle/cvlab.py (Lab conversions, non-local means, the whole enhance_image
restatement) against the installed cv2 and against tests/golden/enhance.npz, which the real
reference's enhance_image (infer_enhanced_burr.py:38-66) produced."""
import os
import zlib

import cv2
import numpy as np
import pytest

import unet_b200.synth as S
from oracle import cvlab, cvnp

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def g():
    return np.load(os.path.join(GOLDEN, "enhance.npz"))


def cube(b_values):
    gg, r = np.meshgrid(np.arange(256, dtype=np.uint8), np.arange(256, dtype=np.uint8), indexing="ij")
    return np.concatenate([np.stack([np.full_like(gg, b), gg, r], -1) for b in b_values], 0)


CODES = {"bgr2lab": (cv2.COLOR_BGR2LAB, cvlab.bgr2lab_u8, True), "lbgr2lab": (cv2.COLOR_LBGR2Lab, cvlab.bgr2lab_u8, False),
         "lab2bgr": (cv2.COLOR_LAB2BGR, cvlab.lab2bgr_u8, True), "lab2lbgr": (cv2.COLOR_Lab2LBGR, cvlab.lab2bgr_u8, False)}


@pytest.mark.parametrize("name", list(CODES))
def test_lab_conversion_vs_cv2(name):
    """Every 8th blue plane (2M colours per code) here; all 2^24 were checked when the oracle
    was written and are checked on the GPU side against the golden checksums."""
    code, fn, srgb = CODES[name]
    img = cube(range(3, 256, 8))
    assert np.array_equal(fn(img, srgb), cv2.cvtColor(img, code))


def test_lab_full_cube_checksum(g):
    """One full 2^24 sweep of the oracle against the checksum cv2 produced in the build container."""
    img = cube(range(256))
    assert zlib.crc32(cvlab.bgr2lab_u8(img, True).tobytes()) == int(g["cube_crc_bgr2lab"])


def test_soft_cbrt_matches_cv2_cuberoot_up_to_truncation():
    x = (np.float32(1.0) / np.float32(2040.0)) * np.arange(1, 3072).astype(np.float32)
    ref = np.array([cv2.cubeRoot(float(v)) for v in x], np.float32)
    got = cvlab.soft_cbrt(x)
    ulp = np.abs(got.view(np.int32) - ref.view(np.int32))
    assert ulp.max() <= 1 and (got <= ref).all()          # truncation never rounds up


def test_nlm_weight_table():
    tab, shift, mult = cvlab.nlm_weight_table(10.0, 1)
    assert (shift, mult, int(tab[0])) == (6, 19096, 19096)
    assert tab[np.nonzero(tab)[0].max()] >= 0.001 * mult
    tab2, _, _ = cvlab.nlm_weight_table(10.0, 2)
    assert np.nonzero(tab2)[0].max() < 2048 and np.nonzero(tab)[0].max() < 2048     # the device table size


@pytest.mark.parametrize("shape", [(40, 56), (33, 47), (14, 90)])
def test_nlm_vs_cv2(shape):
    rng = np.random.default_rng(shape[0])
    base = rng.integers(0, 256, shape + (3,), dtype=np.uint8)
    smooth = cv2.GaussianBlur(base, (9, 9), 3)
    mix = np.clip(smooth.astype(int) + rng.integers(-12, 13, smooth.shape), 0, 255).astype(np.uint8)
    for im in (base, smooth, mix):
        l, ab = np.ascontiguousarray(im[..., 0]), np.ascontiguousarray(im[..., 1:])
        assert np.array_equal(cvlab.nlm_u8(l, 10), cv2.fastNlMeansDenoising(l, None, 10, 7, 21))
        assert np.array_equal(cvlab.nlm_u8(ab, 10), cv2.fastNlMeansDenoising(ab, None, 10, 7, 21))
    assert np.array_equal(cvlab.nlm_colored(mix), cv2.fastNlMeansDenoisingColored(mix, None, 10, 10, 7, 21))


@pytest.mark.parametrize("tag", ["cable", "noise", "speckle"])
def test_enhance_golden(g, tag):
    fr = g[f"{tag}_in"]
    st = cvlab.enhance_stages(fr)
    assert np.array_equal(st["out"], g[f"{tag}_out"])
    lab = st["lab"]
    assert np.array_equal(cvlab.nlm_u8(np.ascontiguousarray(lab[..., 0]), 10), g[f"{tag}_nlm1"])
    assert np.array_equal(cvlab.nlm_u8(np.ascontiguousarray(lab[..., 1:]), 10), g[f"{tag}_nlm2"])


def test_enhance_golden_stage2_fixture(golden_stage2):
    assert np.array_equal(cvlab.enhance(golden_stage2["enh_in"]), golden_stage2["enh_out"])


def test_clahe_sharpen_vs_cv2():
    fr = S.frames_cable(1, 64, 96, seed=9)[0]
    l = np.ascontiguousarray(cv2.cvtColor(fr, cv2.COLOR_BGR2LAB)[..., 0])
    assert np.array_equal(cvnp.clahe_u8(l, 3.0, (8, 8)), cv2.createCLAHE(clipLimit=3.0, tileGridSize=(8, 8)).apply(l))
    k = np.array([[-1, -1, -1], [-1, 9, -1], [-1, -1, -1]])
    assert np.array_equal(cvnp.sharpen3(fr), cv2.filter2D(fr, -1, k))
