"""GPU parity: the vector half of visualize_two_stage on the device (csrc/annotate.cu) against OpenCV, byte for byte.
ub_draw_contours == cv2.findContours(RETR_EXTERNAL, CHAIN_APPROX_SIMPLE) + cv2.drawContours(thickness 2)
(infer_two_stage_burr.py:160-163); ub_paint_layer == drawing a solid-colour shape; all calls through the C ABI."""
import cv2
import numpy as np
import pytest
import torch

import unet_b200.synth as S

pytestmark = pytest.mark.gpu


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def cv_contours(img, mask, colour, thick=2):
    out = img.copy()
    cs, _ = cv2.findContours(np.ascontiguousarray(mask), cv2.RETR_EXTERNAL, cv2.CHAIN_APPROX_SIMPLE)
    cv2.drawContours(out, cs, -1, colour, thick)
    return out


def masks_of_kind(kind, h, w, rng):
    if kind == "noise":
        return (rng.random((h, w)) < rng.uniform(0.05, 0.9)).astype(np.uint8)
    if kind == "blobs":
        return S.mask_blobs(h, w, seed=int(rng.integers(1 << 30)), fill=float(rng.uniform(0.05, 0.6)))
    if kind == "lines":
        m = np.zeros((h, w), np.uint8)
        for _ in range(int(rng.integers(1, 8))):
            cv2.line(m, (int(rng.integers(0, w)), int(rng.integers(0, h))), (int(rng.integers(0, w)), int(rng.integers(0, h))), 255,
                     int(rng.integers(1, 4)))
        return m
    if kind == "nested":            # rings with islands inside: RETR_EXTERNAL keeps the outermost borders only
        m = np.zeros((h, w), np.uint8)
        for _ in range(4):
            c = (int(rng.integers(0, w)), int(rng.integers(0, h)))
            r = int(rng.integers(4, max(5, min(h, w) // 2)))
            cv2.circle(m, c, r, 1, int(rng.integers(1, 4)))
            cv2.circle(m, c, max(r // 3, 1), 1, -1)
        return m
    if kind == "full":
        return np.ones((h, w), np.uint8)
    m = cv2.dilate((rng.random((h, w)) < 0.03).astype(np.uint8), np.ones((3, 3), np.uint8))     # speckle
    m[rng.random((h, w)) < 0.1] = 0
    return m


@pytest.mark.parametrize("hw", [(448, 800), (37, 53), (64, 96), (97, 131), (31, 33), (5, 7), (120, 1024)])
def test_draw_contours_equals_opencv(hw):
    from unet_b200 import _lib
    h, w = hw
    kinds = ["noise", "blobs", "lines", "nested", "speckle", "full"]
    rng = np.random.default_rng(h * 1000 + w)
    masks = np.stack([masks_of_kind(k, h, w, rng) for k in kinds])
    frames = rng.integers(0, 256, (len(kinds), h, w, 3), dtype=np.uint8)
    c = _lib.Context(0, max_batch=len(kinds), frame_hw=hw, net_hw=(16, 16))
    try:
        got = c.draw_contours(dev(frames), dev(masks), (0, 255, 0)).cpu().numpy()
        for i, k in enumerate(kinds):
            want = cv_contours(frames[i], masks[i], (0, 255, 0))
            bad = np.argwhere((got[i] != want).any(axis=2))
            assert len(bad) == 0, f"{k} {hw}: {len(bad)} pixels differ, first at {bad[0].tolist()}"
        # a second call on the same context (the exterior plane is scratch that keeps the last call's contents)
        got2 = c.draw_contours(dev(frames[::-1].copy()), dev(masks[::-1].copy()), (255, 0, 0)).cpu().numpy()
        for i in range(len(kinds)):
            j = len(kinds) - 1 - i
            assert np.array_equal(got2[i], cv_contours(frames[j], masks[j], (255, 0, 0))), kinds[j]
    finally:
        c.close()


@pytest.mark.parametrize("hw", [(448, 800), (64, 96), (37, 53)])
def test_draw_contours_thickness3(hw):
    """Thickness 3 (the burr contours, infer_two_stage_burr.py:165-166): byte-identical to OpenCV when no border pixel of the
    mask lies within a pixel of the image border; otherwise redo is set and OpenCV's result only ADDS pixels of the same
    colour, so redrawing with cv2 on top gives OpenCV's image."""
    from unet_b200 import _lib
    h, w = hw
    kinds = ["noise", "blobs", "lines", "nested", "speckle", "blobs", "lines", "speckle"]
    rng = np.random.default_rng(h + w)
    masks = np.stack([masks_of_kind(k, h, w, rng) for k in kinds])
    masks[4:, :2] = 0; masks[4:, -2:] = 0; masks[4:, :, :2] = 0; masks[4:, :, -2:] = 0      # the last four keep clear of the border
    frames = rng.integers(0, 256, (len(kinds), h, w, 3), dtype=np.uint8)
    col = (255, 0, 255)
    c = _lib.Context(0, max_batch=len(kinds), frame_hw=hw, net_hw=(16, 16))
    try:
        got, redo = c.draw_contours(dev(frames), dev(masks), col, thickness=3)
        got, redo = got.cpu().numpy(), redo.cpu().numpy()
        assert not redo[4:].any(), redo
        for i, k in enumerate(kinds):
            want = cv_contours(frames[i], masks[i], col, 3)
            if not redo[i]:
                assert np.array_equal(got[i], want), (k, i)
            else:
                differs = (got[i] != want).any(axis=2)
                assert (want[differs] == np.array(col, np.uint8)).all(), (k, i)          # only pixels OpenCV paints and the device left
                assert np.array_equal(cv_contours(got[i], masks[i], col, 3), want), (k, i)
    finally:
        c.close()


def test_paint_layer():
    from unet_b200 import _lib
    h, w = 448, 800
    rng = np.random.default_rng(3)
    frames = rng.integers(0, 256, (3, h, w, 3), dtype=np.uint8)
    c = _lib.Context(0, max_batch=3, frame_hw=(h, w), net_hw=(16, 16))
    try:
        layer = np.zeros((h, w), np.uint8)
        cv2.rectangle(layer, (218, 0), (421, 448), 1, 2)
        cv2.putText(layer, "ROI", (223, 25), cv2.FONT_HERSHEY_SIMPLEX, 0.6, 1, 2)
        got = c.paint_layer(dev(frames), dev(layer), [(0, 255, 255)]).cpu().numpy()
        for i in range(3):
            want = frames[i].copy()
            cv2.rectangle(want, (218, 0), (421, 448), (0, 255, 255), 2)
            cv2.putText(want, "ROI", (223, 25), cv2.FONT_HERSHEY_SIMPLEX, 0.6, (0, 255, 255), 2)
            assert np.array_equal(got[i], want)
        per = rng.integers(0, 4, (3, h, w), dtype=np.uint8)
        cols = [(1, 2, 3), (40, 50, 60), (255, 0, 255)]
        got = c.paint_layer(dev(frames), dev(per), cols).cpu().numpy()
        want = frames.copy()
        for k, col in enumerate(cols):
            want[per == k + 1] = col
        assert np.array_equal(got, want)
    finally:
        c.close()


@pytest.mark.parametrize("box", [True, False])
def test_visualize_two_stage_vs_oracle(box):
    """ops.visualize_two_stage (blend, box and the three contour sets on the device; cv2 redraws the burr contours of frames whose
    burr mask touches the border band) == the oracle's restatement of the reference function."""
    import unet_b200 as U
    from oracle import ref_pipeline as R
    rng = np.random.default_rng(11)
    fr = S.frames_cable(2, 448, 800, seed=8)
    roi = R.map_roi((800, 448))
    for i in range(2):
        mc = S.mask_cable_synth(448, 800)
        mt = S.mask_blobs(448, 800, seed=20 + i, fill=0.03)
        mb = S.mask_blobs(448, 800, seed=30 + i, fill=0.004)
        if i == 0:
            mb[:3, 300:330] = 1                              # a burr component on the top border: the redo path
        got = U.visualize_two_stage(fr[i], mc, mt, mb, roi, box)
        want = R.visualize_two_stage(fr[i], mc, mt, mb, roi, box)
        assert got.dtype == np.uint8 and np.array_equal(got, want), int((got != want).any(axis=2).sum())


def test_cli_batch_overlay_path_equals_oracle():
    """What the CLIs do for the result video (cli_common.py): ub_overlay + ops.annotate_on_device on the whole batch, one
    download, then draw_overlay(redo=...) per frame -- the frames equal the oracle's visualize_two_stage."""
    from oracle import ref_pipeline as R
    from unet_b200 import _lib, cli_common, ops
    h, w, n = 448, 800, 3
    fr = S.frames_cable(n, h, w, seed=12)
    roi = R.map_roi((w, h))
    mc = np.stack([S.mask_cable_synth(h, w)] * n)
    mt = np.stack([S.mask_blobs(h, w, seed=40 + i, fill=0.03) for i in range(n)])
    mb = np.stack([S.mask_blobs(h, w, seed=50 + i, fill=0.004) for i in range(n)])
    mb[1, -2:, 250:290] = 1                                   # frame 1: burr pixels on the bottom border -> redo
    c = _lib.Context(0, max_batch=n, frame_hw=(h, w), net_hw=(16, 16))
    try:
        d, dmc, dmt, dmb = dev(fr), dev(mc), dev(mt), dev(mb)
        blend = c.overlay(d, dmc, dmt, dmb, roi)
        redo = ops.annotate_on_device(c, blend, dmc, dmt, dmb, roi).cpu().numpy()
        assert redo[1] == 1
        frames = blend.cpu().numpy()
        for i in range(n):
            got = cli_common.draw_overlay(frames[i], mc[i], mt[i], mb[i], roi, redo=int(redo[i]))
            assert np.array_equal(got, R.visualize_two_stage(fr[i], mc[i], mt[i], mb[i], roi, True)), i
        # frames too large for the device path: cv2 draws everything (redo=None)
        got = cli_common.draw_overlay(c.overlay(d, dmc, dmt, dmb, roi).cpu().numpy()[0], mc[0], mt[0], mb[0], roi, redo=None)
        assert np.array_equal(got, R.visualize_two_stage(fr[0], mc[0], mt[0], mb[0], roi, True))
    finally:
        c.close()


def test_annotate_error_behaviour():
    """Bad arguments to the annotation entry points come back as UbError (negative code + message); the context stays usable."""
    import ctypes as C
    from unet_b200 import _lib
    h, w = 64, 96
    c = _lib.Context(0, max_batch=2, frame_hw=(h, w), net_hw=(16, 16))
    try:
        fr = dev(np.zeros((2, h, w, 3), np.uint8))
        m = dev(np.ones((2, h, w), np.uint8))
        with pytest.raises(_lib.UbError, match="thickness"):
            c.draw_contours(fr, m, (1, 2, 3), thickness=4)
        with pytest.raises(_lib.UbError, match="redo"):                           # thickness 3 needs the flags
            c._check(_lib.lib().ub_draw_contours(c._h, fr.data_ptr(), m.data_ptr(), 2, h, w, 1, 2, 3, 3, None, c.stream))
        with pytest.raises(_lib.UbError, match="exceed the context"):
            c.draw_contours(dev(np.zeros((3, h, w, 3), np.uint8)), dev(np.ones((3, h, w), np.uint8)), (1, 2, 3))
        with pytest.raises(_lib.UbError, match="null"):
            c._check(_lib.lib().ub_paint_layer(c._h, fr.data_ptr(), None, 0, 2, h, w, (C.c_uint8 * 9)(), c.stream))
        out = c.draw_contours(fr, m, (9, 8, 7)).cpu().numpy()                       # a full mask: the frame's outline
        assert (out[0, 0, 0] == (9, 8, 7)).all() and (out[0, h // 2, w // 2] == 0).all()
    finally:
        c.close()
