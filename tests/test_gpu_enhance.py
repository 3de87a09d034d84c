"""GPU parity for enhance_image (infer_enhanced_burr.py:38-66): every step and
the whole function must be bit-exact against the oracle (oracle/cvlab.py, oracle/cvnp.py) and
the golden outputs of the real reference.  All calls go through the C ABI."""
import os
import zlib

import numpy as np
import pytest
import torch

import unet_b200.synth as S
from oracle import cvlab, cvnp

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def g():
    return np.load(os.path.join(GOLDEN, "enhance.npz"))


@pytest.fixture(scope="module")
def ctx():
    from unet_b200 import _lib
    c = _lib.Context(0, max_batch=4, frame_hw=(448, 800), net_hw=(512, 512))
    yield c
    c.close()


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def same(name, got, want):
    got = got.cpu().numpy() if isinstance(got, torch.Tensor) else got
    assert got.shape == want.shape, (name, got.shape, want.shape)
    bad = np.argwhere(got != want)
    assert len(bad) == 0, f"{name}: {len(bad)} of {want.size} differ; first at {bad[0].tolist()} got {got[tuple(bad[0])]} want {want[tuple(bad[0])]}"


def cube(b_values):
    gg, r = np.meshgrid(np.arange(256, dtype=np.uint8), np.arange(256, dtype=np.uint8), indexing="ij")
    return np.concatenate([np.stack([np.full_like(gg, b), gg, r], -1) for b in b_values], 0)


@pytest.mark.parametrize("code,name,srgb", [(0, "bgr2lab", True), (1, "lbgr2lab", False), (2, "lab2bgr", True), (3, "lab2lbgr", False)])
def test_lab_conversion_all_colours(g, code, name, srgb):
    """All 2^24 inputs: checksum from cv2 (golden) for the whole cube, oracle on 32 blue planes."""
    from unet_b200 import _lib
    c = _lib.Context(0, max_batch=1, frame_hw=(256 * 256, 256), net_hw=(16, 16))
    img = cube(range(256))                                  # [65536, 256, 3]
    out = c.cvt_lab(dev(img[None]), code)[0].cpu().numpy()
    c.close()
    assert zlib.crc32(out.tobytes()) == int(g[f"cube_crc_{name}"])
    fn = cvlab.bgr2lab_u8 if code < 2 else cvlab.lab2bgr_u8
    for b in range(5, 256, 8):
        same(f"{name} plane {b}", out[b * 256:(b + 1) * 256], fn(img[b * 256:(b + 1) * 256], srgb))


def test_lab_ragged_batch(ctx):
    rng = np.random.default_rng(3)
    img = rng.integers(0, 256, (3, 37, 53, 3), dtype=np.uint8)       # pixel count not a multiple of 4
    for code, fn, srgb in ((0, cvlab.bgr2lab_u8, True), (3, cvlab.lab2bgr_u8, False)):
        same(f"code {code}", ctx.cvt_lab(dev(img), code), fn(img, srgb))


@pytest.mark.parametrize("shape", [(448, 800), (64, 96), (72, 104)])
def test_clahe(ctx, shape):
    fr = np.concatenate([S.frames_cable(1, *shape, seed=1), S.frames_noise(1, *shape, seed=2)])
    l = np.ascontiguousarray(np.stack([cvlab.bgr2lab_u8(f)[..., 0] for f in fr]))
    want = np.stack([cvnp.clahe_u8(x, 3.0, (8, 8)) for x in l])
    same("clahe", ctx.clahe(dev(l)), want)


def test_clahe_rejects_ragged(ctx):
    from unet_b200 import _lib
    with pytest.raises(_lib.UbError):
        ctx.clahe(dev(np.zeros((1, 70, 100), np.uint8)))


def test_sharpen(ctx):
    rng = np.random.default_rng(5)
    for shape in ((2, 33, 47, 3), (1, 448, 800, 3)):
        img = rng.integers(0, 256, shape, dtype=np.uint8)
        same("sharpen", ctx.sharpen3(dev(img)), np.stack([cvnp.sharpen3(x) for x in img]))


def _nlm_inputs(shape, seed):
    import cv2
    rng = np.random.default_rng(seed)
    base = rng.integers(0, 256, shape + (3,), dtype=np.uint8)
    smooth = cv2.GaussianBlur(base, (9, 9), 3)
    mix = np.clip(smooth.astype(int) + rng.integers(-12, 13, smooth.shape), 0, 255).astype(np.uint8)
    return np.stack([base, smooth, mix])


@pytest.mark.parametrize("shape", [(40, 56), (33, 47), (14, 230), (70, 209)])
def test_nlm_small_and_ragged(ctx, shape):
    ims = _nlm_inputs(shape, shape[1])
    l = np.ascontiguousarray(ims[..., 0])
    ab = np.ascontiguousarray(ims[..., 1:])
    same("nlm 1ch", ctx.nlm(dev(l)), np.stack([cvlab.nlm_u8(x, 10) for x in l]))
    same("nlm 2ch", ctx.nlm(dev(ab)), np.stack([cvlab.nlm_u8(x, 10) for x in ab]))


def test_nlm_rejects_tiny(ctx):
    from unet_b200 import _lib
    with pytest.raises(_lib.UbError):
        ctx.nlm(dev(np.zeros((1, 10, 40), np.uint8)))


@pytest.mark.parametrize("tag", ["cable", "noise", "speckle"])
def test_enhance_golden_small(ctx, g, tag):
    fr = g[f"{tag}_in"]
    same(tag, ctx.enhance(dev(fr[None]))[0], g[f"{tag}_out"])
    lab = cvlab.bgr2lab_u8(fr)
    same(tag + " nlm1", ctx.nlm(dev(np.ascontiguousarray(lab[None, ..., 0])))[0], g[f"{tag}_nlm1"])
    same(tag + " nlm2", ctx.nlm(dev(np.ascontiguousarray(lab[None, ..., 1:])))[0], g[f"{tag}_nlm2"])


def test_enhance_golden_stage2_fixture(ctx, golden_stage2):
    same("enh 128x256", ctx.enhance(dev(golden_stage2["enh_in"][None]))[0], golden_stage2["enh_out"])


def test_enhance_full_frame_vs_oracle_stages(ctx, g):
    """800x448: every intermediate of the device path against the oracle, the result against the
    checksum / crop of the real reference's output."""
    fr = S.frames_cable(1, 448, 800, seed=0)[0]
    assert zlib.crc32(fr.tobytes()) == int(g["full_in_crc"])
    st = cvlab.enhance_stages(fr)
    d = dev(fr[None])
    lab = ctx.cvt_lab(d, 0)
    same("lab", lab[0], st["lab"])
    l2 = ctx.clahe(lab[..., 0].contiguous())
    same("clahe", l2[0], st["l_clahe"])
    lab2 = torch.cat([l2[..., None], lab[..., 1:]], -1).contiguous()
    bgr = ctx.cvt_lab(lab2, 2)
    same("bgr_clahe", bgr[0], st["bgr_clahe"])
    lin = ctx.cvt_lab(bgr, 1)
    same("lab_lin", lin[0], st["lab_lin"])
    same("l_den", ctx.nlm(lin[..., 0].contiguous())[0], st["l_den"])
    same("ab_den", ctx.nlm(lin[..., 1:].contiguous())[0], st["ab_den"])
    out = ctx.enhance(d)[0].cpu().numpy()
    same("out", out, st["out"])
    assert zlib.crc32(out.tobytes()) == int(g["full_out_crc"])
    same("crop", out[192:256, 368:432], g["full_out_crop"])


def test_enhance_batch_and_ops_contract(ctx):
    """Batch of 4 equals 4 single calls; the reference-facing ops.enhance_image takes and returns numpy."""
    import cv2
    import unet_b200 as U
    fr = np.concatenate([S.frames_cable(2, 448, 800, seed=11), S.frames_noise(1, 448, 800, seed=12),
                         S.frames_speckle(1, 448, 800, seed=13)])
    out = ctx.enhance(dev(fr)).cpu().numpy()
    k = np.array([[-1, -1, -1], [-1, 9, -1], [-1, -1, -1]])
    for i in range(4):
        lab = cv2.cvtColor(fr[i], cv2.COLOR_BGR2LAB)
        lab[..., 0] = cv2.createCLAHE(clipLimit=3.0, tileGridSize=(8, 8)).apply(np.ascontiguousarray(lab[..., 0]))
        want = cv2.filter2D(cv2.fastNlMeansDenoisingColored(cv2.cvtColor(lab, cv2.COLOR_LAB2BGR), None, 10, 10, 7, 21), -1, k)
        same(f"frame {i}", out[i], want)
    one = U.enhance_image(fr[0])
    assert isinstance(one, np.ndarray) and one.dtype == np.uint8
    same("ops", one, out[0])


def test_pipeline_with_enhance_flag(golden_net):
    """--enhance: the enhanced frame feeds the net AND the gray image (infer_enhanced_burr.py:279-283)."""
    from unet_b200 import InspectionPipeline
    sd = S.make_state_dict(seed=0)
    fr = S.frames_cable(2, 448, 800, seed=21)
    pe = InspectionPipeline(sd, batch=2, variant="enhanced", enhance=True)
    pn = InspectionPipeline(sd, batch=2, variant="enhanced", enhance=False)
    se, me = pe.run(dev(fr), want_masks=True)
    enh = pe.ctx.enhance(dev(fr))
    sn, mn = pn.run(enh, want_masks=True)
    same("stats", se, sn.cpu().numpy())
    same("masks", me, mn.cpu().numpy())
    host = pe.run_host(torch.from_numpy(fr).pin_memory())
    same("host stats", host, sn.cpu().numpy())
