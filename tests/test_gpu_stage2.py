"""GPU parity: integer path (pre/post-processing and the burr stage) must be bit-exact against
the oracle (cv2-backed restatement in oracle/ref_pipeline.py + numpy spec in oracle/cvnp.py).
All calls go through the C ABI (unet_b200._lib -> libunetb200.so)."""
import cv2
import numpy as np
import pytest
import torch

import unet_b200.synth as S
from oracle import cvnp
from oracle import ref_pipeline as R

pytestmark = pytest.mark.gpu


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


def frames4():
    return np.concatenate([S.frames_cable(2, 448, 800, seed=1), S.frames_noise(1, 448, 800, seed=2),
                           S.frames_speckle(1, 448, 800, seed=5)])


def test_preprocess_and_gray(ctx):
    fr = frames4()
    x, g = ctx.preprocess(dev(fr))
    for i in range(4):
        same(f"gray[{i}]", g[i], cv2.cvtColor(fr[i], cv2.COLOR_BGR2GRAY))
        same(f"x[{i}]", x[i], R.preprocess(fr[i]).numpy())


def test_preprocess_golden(ctx, golden_stage2, golden_frames):
    for tag in ("cable", "noise"):
        fr = golden_frames[tag]
        c = ctx
        if fr.shape[:2] != (448, 800):
            from unet_b200 import _lib
            c = _lib.Context(0, 1, fr.shape[:2], (512, 512))
        x, g = c.preprocess(dev(fr[None]))
        same(tag + " gray", g[0], golden_stage2[f"{tag}_gray"])
        same(tag + " pre", x[0], golden_stage2[f"{tag}_pre_u8"].astype(np.float32) / 255.0)


@pytest.mark.parametrize("hw,net", [((301, 397), (256, 256)), ((1080, 1920), (512, 512)), ((97, 131), (512, 512)), ((448, 800), (512, 512))])
def test_preprocess_sizes_and_alignment(hw, net):
    """preprocess_image at sizes whose rows are not 16-byte multiples (shared-memory staging starts mid-vector), a
    downscale by more than 2 (no row shared between vertical taps), a strong upscale, and a frame pointer 1 byte off
    alignment (first / last staged vector falls outside the buffer)."""
    from unet_b200 import _lib
    H, W = hw
    c = _lib.Context(0, 3, hw, net)
    try:
        fr = S.frames_noise(3, H, W, seed=H + W)
        x, _ = c.preprocess(dev(fr), want_gray=False)
        for i in range(3):
            same(f"x[{i}]", x[i], R.preprocess(fr[i], size=(net[1], net[0])).numpy())
        raw = torch.zeros(fr.size + 1, dtype=torch.uint8, device="cuda")
        raw[1:] = dev(fr).reshape(-1)
        x, _ = c.preprocess(raw[1:].view(3, H, W, 3), want_gray=False)
        same("x unaligned", x[2], R.preprocess(fr[2], size=(net[1], net[0])).numpy())
    finally:
        c.close()


def test_normalize_frames(ctx, golden_stage2):
    big = S.frames_noise(2, 1080, 1920, seed=3)
    out = ctx.normalize_frames(dev(big), (800, 448), rotate=False)
    for i in range(2):
        same("resize", out[i], cv2.resize(big[i], (800, 448), interpolation=cv2.INTER_LINEAR))
    out = ctx.normalize_frames(dev(big), (800, 448), rotate=True)
    for i in range(2):
        same("rot+resize", out[i], R.normalize_frame(big[i], rotate=True))
    g = golden_stage2
    out = ctx.normalize_frames(dev(g["norm_in"][None]), (800, 448), rotate=True)
    same("golden normalize", out[0], g["norm_out"])


def test_masks_from_logits(ctx):
    rng = np.random.default_rng(0)
    logits = rng.normal(size=(2, 3, 512, 512)).astype(np.float32)
    logits[0, :, :40] = 0.25                      # exact ties -> lowest index
    roi = R.map_roi((800, 448))
    pred, cable, tape = ctx.masks_from_logits(dev(logits), (448, 800), roi)
    for i in range(2):
        p, c, t = R.masks_from_logits(torch.from_numpy(logits[i]), (800, 448), roi)
        same("pred", pred[i], p)
        same("cable", cable[i], c)
        same("tape", tape[i], t)


@pytest.mark.parametrize("hw", [(448, 800), (37, 53), (64, 96), (36, 164), (100, 324), (33, 160), (31, 4), (70, 484)])
def test_gauss_canny_edges(ctx, hw):
    h, w = hw
    fr = np.concatenate([S.frames_cable(1, h, w, seed=1), S.frames_noise(1, h, w, seed=2)])
    gray = np.stack([cv2.cvtColor(f, cv2.COLOR_BGR2GRAY) for f in fr])
    blur = ctx.gaussian5(dev(gray))
    ref_blur = np.stack([cv2.GaussianBlur(g, (5, 5), 1.0) for g in gray])
    same("gauss", blur, ref_blur)
    for lo, hi in ((50, 150), (30, 100)):
        e = ctx.canny(dev(ref_blur), lo, hi)
        same(f"canny{lo}", e, np.stack([cv2.Canny(b, lo, hi) for b in ref_blur]))
        same(f"blur_canny{lo}", ctx.blur_canny(dev(gray), lo, hi), np.stack([cv2.Canny(b, lo, hi) for b in ref_blur]))
    same("sobel", ctx.sobel_edges(dev(gray)), np.stack([cvnp.sobel_edges(g) for g in gray]))
    same("lap", ctx.laplacian_edges(dev(gray)), np.stack([cvnp.laplacian_edges(g) for g in gray]))


@pytest.mark.parametrize("k", [2, 3, 5, 8, 25])
def test_morphology(ctx, k):
    se = cv2.getStructuringElement(cv2.MORPH_ELLIPSE, (k, k))
    for h, w, fill in ((448, 800, 0.3), (50, 70, 0.05), (31, 33, 0.6), (448, 800, 0.002)):
        b = np.stack([S.mask_blobs(h, w, seed=k + i, fill=fill) for i in range(2)])
        d = dev(b)
        same(f"dilate{k}", ctx.morph("dilate", k, d), np.stack([cv2.dilate(x, se) for x in b]))
        same(f"erode{k}", ctx.morph("erode", k, d), np.stack([cv2.erode(x, se) for x in b]))
        same(f"close{k}", ctx.morph("close", k, d), np.stack([cv2.morphologyEx(x, cv2.MORPH_CLOSE, se) for x in b]))
        same(f"open{k}", ctx.morph("open", k, d), np.stack([cv2.morphologyEx(x, cv2.MORPH_OPEN, se) for x in b]))


@pytest.mark.parametrize("hw,fill", [((448, 800), 0.3), ((447, 799), 0.45), ((33, 65), 0.6), ((448, 800), 0.02)])
def test_ccl_labels_and_stats(ctx, hw, fill):
    b = np.stack([S.mask_blobs(hw[0], hw[1], seed=9 + i, fill=fill) for i in range(2)])
    n, lab, st = ctx.ccl(dev(b), max_labels=32768)
    n, lab, st = n.cpu().numpy(), lab.cpu().numpy(), st.cpu().numpy()
    for i in range(2):
        rn, rl, rs, _ = cv2.connectedComponentsWithStats(b[i], connectivity=8)
        assert n[i] == rn
        same("labels", lab[i], rl)
        same("stats", st[i, :rn], rs)


@pytest.mark.parametrize("variant", ["standard", "enhanced"])
def test_burr_stage_golden(ctx, variant, golden_stage2, golden_frames):
    from unet_b200 import _lib
    g = golden_stage2
    for tag in ("cable", "speckle", "dots", "noise"):
        fr = golden_frames[tag]
        gray = cv2.cvtColor(fr, cv2.COLOR_BGR2GRAY)
        mask = g[f"{tag}_mask"]
        c = ctx if gray.shape == (448, 800) else _lib.Context(0, 1, gray.shape, (512, 512))
        if variant == "standard":
            for sens in ("medium", "high"):
                cfg = R.BURR_CONFIGS[sens]
                out, px = c.burr(dev(gray[None]), dev(mask[None]), 0, _lib.standard_params(cfg["min_area"], cfg["max_area"]))
                same(f"{tag} burr {sens}", out[0], g[f"{tag}_burr_{sens}"])
                assert int(px[0]) == int(g[f"{tag}_burr_{sens}"].sum())
            out, px = c.burr(dev(gray[None]), dev(np.zeros_like(mask)[None]), 0)
            same(f"{tag} empty", out[0], g[f"{tag}_burr_empty"])
        else:
            out, px = c.burr(dev(gray[None]), dev(mask[None]), 1)
            same(f"{tag} burr enh", out[0], g[f"{tag}_burr_enh"])
            assert int(px[0]) == int(g[f"{tag}_burr_enh"].sum())


def test_burr_stage_batched_vs_oracle(ctx):
    fr = frames4()
    gray = np.stack([cv2.cvtColor(f, cv2.COLOR_BGR2GRAY) for f in fr])
    masks = np.stack([S.mask_cable_synth(448, 800), S.mask_blobs(448, 800, 3, 0.002),
                      S.mask_cable_synth(448, 800, R.map_roi((800, 448))), S.mask_blobs(448, 800, 4, 0.01)])
    out, px = ctx.burr(dev(gray), dev(masks), 0)
    oute, pxe = ctx.burr(dev(gray), dev(masks), 1)
    for i in range(4):
        same(f"std[{i}]", out[i], R.burrs_standard(gray[i], masks[i]))
        same(f"enh[{i}]", oute[i], R.burrs_enhanced(gray[i], masks[i]))
    assert px.cpu().tolist() == [int(out[i].sum()) for i in range(4)]


def test_function_contract():
    """The reference-named functions accept numpy like the reference and return numpy."""
    import unet_b200 as U
    fr = S.frames_speckle(1, 448, 800, seed=5)[0]
    gray = cv2.cvtColor(fr, cv2.COLOR_BGR2GRAY)
    mask = S.mask_blobs(448, 800, seed=3, fill=0.002)
    same("detect_burrs_on_cable", U.detect_burrs_on_cable(gray, mask, {"min_area": 20, "max_area": 1000}),
         R.burrs_standard(gray, mask, {"min_area": 20, "max_area": 1000}))
    same("detect_burrs_enhanced", U.detect_burrs_enhanced(gray, mask, {"min_area": 50, "max_area": 500}),
         R.burrs_enhanced(gray, mask))
    x = U.preprocess_image(fr)
    assert isinstance(x, torch.Tensor) and not x.is_cuda and x.dtype == torch.float32
    same("preprocess_image", x.numpy(), R.preprocess(fr).numpy())
    assert U.map_roi_to_original((800, 448)) == (218, 0, 421, 448)
    same("bgr2gray", U.bgr2gray(fr), gray)


@pytest.mark.parametrize("hw", [(448, 800), (90, 158), (33, 67)])
def test_overlay_blend_vs_oracle(ctx, hw):
    """ub_overlay == the four cv2.addWeighted steps of visualize_two_stage (oracle: cvnp.overlay_blend, which
    tests/test_oracle_cvnp.py pins against cv2 on every operand pair)."""
    h, w = hw
    fr = np.concatenate([S.frames_cable(1, h, w, seed=3), S.frames_noise(2, h, w, seed=4)])
    roi = R.map_roi((w, h))
    mc = np.stack([S.mask_cable_synth(h, w, roi)] * 3)
    mt = np.stack([S.mask_blobs(h, w, seed=8 + i, fill=0.3) for i in range(3)])
    mb = np.stack([S.mask_blobs(h, w, seed=20 + i, fill=0.1) * 255 for i in range(3)]).astype(np.uint8)
    got = ctx.overlay(dev(fr), dev(mc), dev(mt), dev(mb), roi)
    want = np.stack([cvnp.overlay_blend(fr[i], mc[i], mt[i], mb[i], roi) for i in range(3)])
    same("overlay", got, want)
    for i in range(3):       # and the oracle is cv2 itself
        dimmed = fr[i].copy(); outside = np.ones((h, w), bool); outside[roi[1]:roi[3], roi[0]:roi[2]] = False; dimmed[outside] = 0
        ref = cv2.addWeighted(fr[i], 0.7, dimmed, 0.3, 0)
        for m, col, a, b in ((mc[i], (0, 255, 0), .6, .4), (mt[i], (255, 0, 0), .6, .4), (mb[i], (255, 0, 255), .5, .5)):
            tint = np.zeros_like(ref); tint[m > 0] = col
            ref = cv2.addWeighted(ref, a, tint, b, 0)
        same("overlay_cv2", got[i], ref)


@pytest.mark.parametrize("tag", ["full", "small"])
def test_visualize_two_stage_golden(golden_viz, tag):
    """ops.visualize_two_stage (device blend + cv2 drawing) == the real reference's output, byte for byte."""
    import zlib
    from conftest import viz_inputs
    from unet_b200 import ops as U
    fr, mc, mt, mb, roi = viz_inputs(tag)
    box = U.visualize_two_stage(fr, mc, mt, mb, roi, True)
    nobox = U.visualize_two_stage(fr, mc, mt, mb, roi, False)
    assert zlib.crc32(box.tobytes()) == int(golden_viz[f"{tag}_box_crc"])
    assert zlib.crc32(nobox.tobytes()) == int(golden_viz[f"{tag}_nobox_crc"])


def test_rulebased_detector_golden_and_oracle(golden_rulebased):
    """ops.get_burr_mask_rulebased (ub_burr, UB_VARIANT_LAPLACIAN) == the real reference's output (golden) and the oracle
    on other masks / thresholds."""
    from conftest import rulebased_cases, RULEBASED_CFGS
    from unet_b200 import ops as U
    for tag, gray, mask in rulebased_cases():
        for name, cfg in RULEBASED_CFGS.items():
            out = U.get_burr_mask_rulebased(gray, mask, cfg)
            assert out.dtype == np.uint8 and set(np.unique(out)) <= {0, 255}
            assert np.array_equal(np.packbits(out > 0), golden_rulebased[f"{tag}_{name}"]), (tag, name)
        assert U.get_burr_mask_rulebased(gray, np.zeros_like(mask)).max() == 0
        h, w = gray.shape
        blobs = (S.mask_blobs(h, w, seed=11, fill=0.0008) * 255).astype(np.uint8)          # {0,255} mask, many small seeds
        for cfg in (dict(band_out=3, laplacian_threshold=10, min_area=1, max_area=100000), dict(band_out=7, laplacian_threshold=60, min_area=3, max_area=40)):
            same(f"rulebased_{tag}", U.get_burr_mask_rulebased(gray, blobs, cfg), R.burrs_rulebased(gray, blobs, **cfg))


def test_diameter_metrics_golden_and_oracle(golden_geometry):
    """ops.compute_diameter_metrics (ub_diameter_metrics) == the real reference's numbers, every field exactly; batched
    device call == oracle on other class maps, incl. two equal-area components (OpenCV label order breaks the tie)."""
    from conftest import geometry_cases, GEOM_FIELDS
    from unet_b200 import ops as U
    for pred, want, want_short in geometry_cases(golden_geometry):
        m = U.compute_diameter_metrics(pred)
        assert [float(getattr(m, f)) for f in GEOM_FIELDS] == list(want), (pred.shape, m)
        m = U.compute_diameter_metrics(pred, mm_per_px=0.1, min_valid_rows=1000)
        assert [float(getattr(m, f)) for f in GEOM_FIELDS] == list(want_short), (pred.shape, m)
    preds = np.stack([S.pred_classes(448, 800, seed=10 + i, tape=i != 2) for i in range(4)])
    preds[3][:] = 0                                   # tie: two cable blocks of equal area; tape only beside the second
    preds[3][300:360, 100:200] = 1
    preds[3][20:80, 500:600] = 1
    preds[3][295:365, 60:100] = 2
    preds[3][10:90, 460:500] = 2
    preds[3][10:90, 330:400] = 2
    got = U.compute_diameter_metrics(dev(preds))
    for i in range(4):
        want = R.diameter_metrics(preds[i])
        assert [float(getattr(got[i], f)) for f in GEOM_FIELDS] == [float(want[f]) for f in GEOM_FIELDS], (i, got[i], want)
    empty = U.compute_diameter_metrics(np.zeros((64, 64), np.uint8))
    assert empty.valid_rows == 0 and empty.dc_px == 0.0 and empty.cable_coverage == 0.0


def test_analyze_defects_golden_and_oracle(golden_geometry):
    """ops.analyze_defects (ub_analyze_defects) == the real reference's fields exactly; batched device call == oracle."""
    from conftest import defect_cases, defect_row
    from unet_b200 import ops as U
    for pred, hole_min, want in defect_cases(golden_geometry):
        assert defect_row(U.analyze_defects(pred, hole_min_size=hole_min)) == list(want), (pred.shape, hole_min)
    preds = np.stack([S.pred_classes(448, 800, seed=30 + i, tape=i != 1, defects=True) for i in range(3)])
    got = U.analyze_defects(dev(preds), hole_min_size=4)
    for i in range(3):
        assert defect_row(got[i]) == defect_row(R.analyze_defects(preds[i], hole_min_size=4)), i
    z = U.analyze_defects(np.zeros((48, 64), np.uint8))
    assert z.tape_num_components == 0 and z.tape_largest_area_ratio == 0.0 and z.tape_hole_ratio == 0.0 and z.total_defect_area == 0


def test_thickness_profile_golden_and_oracle(golden_geometry):
    """ops.compute_thickness_profile (ub_width_profiles) == the real reference's arrays, bit for bit."""
    from unet_b200 import ops as U
    g = golden_geometry
    for i, (h, w, seed, tape) in enumerate(g["cases"]):
        pred = S.pred_classes(int(h), int(w), seed=int(seed), tape=bool(tape), defects=True)
        t = U.compute_thickness_profile(pred, mm_per_px=0.07)
        assert t.delta_d_mm.dtype == np.float32 and np.array_equal(t.delta_d_mm, g[f"profile_{i}"]), i
        assert np.array_equal(np.packbits(t.valid_mask), g[f"profile_valid_{i}"]) and np.array_equal(t.y_coords, np.arange(h))
    preds = np.stack([S.pred_classes(448, 800, seed=40 + i, tape=i != 1, defects=True) for i in range(3)])
    got = U.compute_thickness_profile(dev(preds))
    for i in range(3):
        d, valid = R.thickness_profile(preds[i])
        assert np.array_equal(got[i].delta_d_mm, d) and np.array_equal(got[i].valid_mask, valid), i


def test_postprocess_filters_golden_and_oracle(golden_geometry):
    """ops.filter_cable_by_shape / ops.constrain_tape_to_ring == the real reference's masks; batched call == oracle."""
    from conftest import postprocess_cases
    from unet_b200 import ops as U
    g = golden_geometry
    for i, name, cfg, cable, tape, w in postprocess_cases(g):
        fc = U.filter_cable_by_shape(cable, cfg, w)
        assert set(np.unique(fc)) <= {0, 255}
        assert np.array_equal(np.packbits(fc > 0), g[f"cable_{name}_{i}"]), (i, name)
        ring = U.constrain_tape_to_ring(tape, cable, cfg)
        assert set(np.unique(ring)) <= {0, 255} and np.array_equal(np.packbits(ring > 0), g[f"ring_{name}_{i}"]), (i, name)
        assert np.array_equal(np.packbits(U.constrain_tape_to_ring(tape * 255, fc, cfg) > 0), g[f"ring255_{name}_{i}"]), (i, name)
    preds = np.stack([S.pred_classes(448, 800, seed=50 + i, tape=True, defects=True) for i in range(3)])
    cable, tape = (preds == 1).astype(np.uint8), (preds == 2).astype(np.uint8)
    cable[2][:] = 0                                      # two identical blocks: equal score, the first label wins
    cable[2][300:400, 380:420] = 1
    cable[2][100:200, 380:420] = 1
    got = U.filter_cable_by_shape(dev(cable), dict(cable_min_area=100, cable_min_aspect=1.2, cable_max_center_offset=0.4), 800)
    ring = U.constrain_tape_to_ring(dev(tape), dev(cable))
    for i in range(3):
        same("filter_cable", got[i], R.filter_cable_by_shape(cable[i], 100, 1.2, 0.4, 800))
        same("tape_ring", ring[i], R.constrain_tape_to_ring(tape[i], cable[i]))
    z = np.zeros((64, 96), np.uint8)
    assert U.filter_cable_by_shape(z).max() == 0 and U.constrain_tape_to_ring(z, z).max() == 0


def test_error_behaviour_through_the_c_abi(ctx):
    """Bad arguments come back as a negative code + message (UbError), never as a crash, and leave the context usable."""
    from unet_b200 import _lib
    g = dev(S.frames_noise(1, 448, 800, seed=1)[..., 0].copy())
    m = dev(S.mask_cable_synth(448, 800)[None])
    with pytest.raises(_lib.UbError, match="exceed the context"):
        ctx.gaussian5(dev(np.zeros((5, 448, 800), np.uint8)))              # batch 5 > max_batch 4
    with pytest.raises(_lib.UbError, match="exceed the context"):
        ctx.canny(dev(np.zeros((1, 900, 800), np.uint8)), 50, 150)         # frame larger than the context's
    with pytest.raises(_lib.UbError, match="ellipse size"):
        ctx.morph("dilate", 27, m)
    with pytest.raises(_lib.UbError, match="variant"):
        ctx.burr(g, m, variant=7, params=_lib.standard_params())
    with pytest.raises(_lib.UbError, match="smaller than 16x16"):
        ctx.diameter_metrics(dev(np.zeros((1, 8, 8), np.uint8)))
    with pytest.raises(_lib.UbError, match="roi_width"):
        ctx.filter_cable_by_shape(m, 10, 1.0, 0.5, 0)
    out, px = ctx.burr(g, m)                                                # still works afterwards
    assert out.shape == g.shape and int(px[0]) == int(out.sum())
