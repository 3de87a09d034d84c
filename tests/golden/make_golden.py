"""Generate tests/golden/*.npz by running the REAL reference (imported from
/root/reference) on seeded synthetic inputs.  Run in the build container only:

    python tests/golden/make_golden.py

The reference ships no tests, fixtures or golden vectors of its own,
so these files are what pins oracle/ to the reference.  Library versions are recorded
inside each file.
"""
import os
import sys

import cv2
import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")

import infer_two_stage_burr as T          # noqa: E402  (the reference)
import infer_enhanced_burr as E           # noqa: E402
from src.models.unetpp import NestedUNet  # noqa: E402

import unet_b200.synth as S               # noqa: E402

cv2.setNumThreads(1)
torch.set_num_threads(4)
VERS = np.array([cv2.__version__, torch.__version__, np.__version__])


def ref_model(sd):
    m = NestedUNet(num_classes=3, deep_supervision=True, pretrained_encoder=False)
    m.load_state_dict(sd, strict=True)
    return m.eval()


def stage2():
    out = {"versions": VERS}
    cable = S.frames_cable(1, 448, 800, seed=0)[0]
    noise = S.frames_noise(1, 160, 256, seed=1)[0]
    speckle = S.frames_speckle(1, 448, 800, seed=5, n_blobs=400)[0]
    for tag, fr in (("cable", cable), ("noise", noise), ("speckle", speckle), ("dots", speckle)):
        h, w = fr.shape[:2]
        gray = cv2.cvtColor(fr, cv2.COLOR_BGR2GRAY)
        roi = T.map_roi_to_original((w, h), (512, 512))
        mask = S.mask_cable_synth(h, w, None)
        if tag == "dots":
            mask = S.mask_blobs(h, w, seed=3, fill=0.002)
        # large frames are regenerated from unet_b200.synth in the tests; crc32 pins them
        import zlib
        out[f"{tag}_crc"] = np.array(zlib.crc32(fr.tobytes()), np.uint32)
        if tag == "noise":
            out[f"{tag}_frame"] = fr
        if tag in ("cable", "noise"):
            out[f"{tag}_gray"] = gray
        out[f"{tag}_roi"] = np.array(roi, np.int32)
        out[f"{tag}_mask"] = mask
        pre = T.preprocess_image(fr).numpy()               # [3,512,512] f32 = u8/255 exactly
        pre_u8 = np.rint(pre * 255.0).astype(np.uint8)
        assert np.array_equal(pre_u8.astype(np.float32) / 255.0, pre)
        if tag in ("cable", "noise"):
            out[f"{tag}_pre_u8"] = pre_u8                  # stored as u8 (4x smaller)
        for sens, cfg in (("medium", {"min_area": 30, "max_area": 800}),
                          ("high", {"min_area": 20, "max_area": 1000})):
            out[f"{tag}_burr_{sens}"] = T.detect_burrs_on_cable(gray, mask, cfg)
        out[f"{tag}_burr_enh"] = E.detect_burrs_enhanced(gray, mask, {"min_area": 50, "max_area": 500})
        out[f"{tag}_burr_empty"] = T.detect_burrs_on_cable(gray, np.zeros_like(mask), None)
    # enhance_image on a small crop-sized frame (NLM is slow) -- size divisible by the 8x8 CLAHE grid
    small = S.frames_cable(1, 128, 256, seed=2)[0]
    out["enh_in"] = small
    out["enh_out"] = E.enhance_image(small)
    big = S.frames_cable(1, 270, 480, seed=3)[0]
    out["norm_in"] = big
    out["norm_out"] = cv2.resize(cv2.rotate(big, cv2.ROTATE_90_COUNTERCLOCKWISE), (800, 448),
                                 interpolation=cv2.INTER_LINEAR)
    np.savez_compressed(os.path.join(HERE, "stage2.npz"), **out)


def net():
    out = {"versions": VERS}
    sd = S.make_state_dict(seed=0)
    m = ref_model(sd)
    keys = list(m.state_dict().keys())
    out["state_dict_keys"] = np.array(keys)
    g = torch.Generator().manual_seed(7)
    for tag, shape in (("a", (1, 3, 64, 64)), ("b", (2, 3, 48, 80))):
        x = torch.rand(*shape, generator=g)
        with torch.no_grad():
            y = m(x)
        out[f"x_{tag}"] = x.numpy()
        out[f"y_{tag}"] = y.numpy()
    # whole frame step through the reference's own statements (infer_two_stage_burr.py:286-323)
    fr = S.frames_cable(1, 448, 800, seed=4)[0]
    roi = T.map_roi_to_original((800, 448), (512, 512))
    x1, y1, x2, y2 = roi
    img = T.preprocess_image(fr).unsqueeze(0)
    with torch.no_grad():
        o = m(img)
    probs = torch.softmax(o, dim=1)[0].cpu().numpy()
    pred = np.argmax(probs, axis=0).astype(np.uint8)
    res = {}
    for name, cls in (("cable", 1), ("tape", 2)):
        full = cv2.resize((pred == cls).astype(np.uint8), (800, 448), interpolation=cv2.INTER_NEAREST)
        mk = np.zeros_like(full)
        mk[y1:y2, x1:x2] = full[y1:y2, x1:x2]
        res[name] = mk
    gray = cv2.cvtColor(fr, cv2.COLOR_BGR2GRAY)
    burr = T.detect_burrs_on_cable(gray, res["cable"], {"min_area": 30, "max_area": 800})
    out["step_seed"] = np.array(4)
    out["step_logits_sub"] = o[0, :, ::8, ::8].numpy()
    out["step_logits_sum"] = np.array([float(o.double().sum()), float(o.double().abs().sum())])
    out["step_pred"] = pred
    out["step_mask_cable"] = res["cable"]
    out["step_mask_tape"] = res["tape"]
    out["step_mask_burr"] = burr
    out["step_counts"] = np.array([res["cable"].sum(), res["tape"].sum(), burr.sum()], np.int64)
    np.savez_compressed(os.path.join(HERE, "net.npz"), **out)


def colour_cube(b_values=None):
    """All 2^24 (or the given blue planes of) uint8 triples as [n*256, 256, 3] images."""
    g, r = np.meshgrid(np.arange(256, dtype=np.uint8), np.arange(256, dtype=np.uint8), indexing="ij")
    bs = range(256) if b_values is None else b_values
    return np.concatenate([np.stack([np.full_like(g, b), g, r], -1) for b in bs], 0)


def enhance():
    """enhance_image (infer_enhanced_burr.py:38-66) of the real reference + the cv2 calls inside it."""
    import zlib
    out = {"versions": VERS}
    cube = colour_cube()
    for name, code in (("bgr2lab", cv2.COLOR_BGR2LAB), ("lbgr2lab", cv2.COLOR_LBGR2Lab),
                       ("lab2bgr", cv2.COLOR_LAB2BGR), ("lab2lbgr", cv2.COLOR_Lab2LBGR)):
        out[f"cube_crc_{name}"] = np.array(zlib.crc32(cv2.cvtColor(cube, code).tobytes()), np.uint32)
    for tag, fr in (("cable", S.frames_cable(1, 64, 96, seed=2)[0]), ("noise", S.frames_noise(1, 72, 104, seed=3)[0]),
                    ("speckle", S.frames_speckle(1, 48, 56, seed=4, n_blobs=12)[0])):
        out[f"{tag}_in"] = fr
        out[f"{tag}_out"] = E.enhance_image(fr)
        lab = cv2.cvtColor(fr, cv2.COLOR_BGR2LAB)
        out[f"{tag}_nlm1"] = cv2.fastNlMeansDenoising(np.ascontiguousarray(lab[..., 0]), None, 10, 7, 21)
        out[f"{tag}_nlm2"] = cv2.fastNlMeansDenoising(np.ascontiguousarray(lab[..., 1:]), None, 10, 7, 21)
    full = S.frames_cable(1, 448, 800, seed=0)[0]
    out["full_in_crc"] = np.array(zlib.crc32(full.tobytes()), np.uint32)
    eo = E.enhance_image(full)
    out["full_out_crc"] = np.array(zlib.crc32(eo.tobytes()), np.uint32)
    out["full_out_crop"] = eo[192:256, 368:432].copy()
    np.savez_compressed(os.path.join(HERE, "enhance.npz"), **out)


def rulebased():
    """get_burr_mask_rulebased of the refactored library (src/refactor/burr_detector.py:11-66)."""
    from src.refactor.burr_detector import get_burr_mask_rulebased
    from src.refactor.config import BurrConfig
    out = {"versions": VERS}
    for tag, fr, mask in (("speckle", S.frames_speckle(1, 448, 800, seed=5, n_blobs=400)[0], S.mask_cable_synth(448, 800, None)),
                          ("noise", S.frames_noise(1, 120, 184, seed=1)[0], S.mask_cable_synth(120, 184, None))):
        gray = cv2.cvtColor(fr, cv2.COLOR_BGR2GRAY)
        for name, cfg in (("default", BurrConfig()), ("wide", BurrConfig(band_out=12, laplacian_threshold=20, min_area=5, max_area=2000))):
            out[f"{tag}_{name}"] = np.packbits(get_burr_mask_rulebased(gray, mask, cfg) > 0)
            out[f"{tag}_{name}_vals"] = np.unique(get_burr_mask_rulebased(gray, mask, cfg))
        out[f"{tag}_empty"] = get_burr_mask_rulebased(gray, np.zeros_like(mask), BurrConfig()).max()
    np.savez_compressed(os.path.join(HERE, "rulebased.npz"), **out)


GEOM_CASES = (((512, 512), 0, True), ((448, 800), 1, True), ((512, 512), 2, False), ((96, 160), 3, True), ((130, 75), 4, True))
GEOM_FIELDS = ("dc_px", "dt_px", "delta_d_px", "dc_mm", "dt_mm", "delta_d_mm", "valid_rows", "cable_coverage", "tape_coverage")


def geometry():
    """compute_diameter_metrics of the real reference (src/utils/geometry_enhanced.py:113-187) on unet_b200.synth.pred_classes."""
    import zlib
    from src.utils.geometry_enhanced import compute_diameter_metrics
    out = {"versions": VERS, "cases": np.array([[h, w, seed, int(t)] for (h, w), seed, t in GEOM_CASES])}
    rows, crcs = [], []
    for (h, w), seed, tape in GEOM_CASES:
        p = S.pred_classes(h, w, seed=seed, tape=tape)
        crcs.append(zlib.crc32(p.tobytes()))
        m = compute_diameter_metrics(p)
        rows.append([float(getattr(m, f)) for f in GEOM_FIELDS])
        m2 = compute_diameter_metrics(p, mm_per_px=0.1, min_valid_rows=1000)       # the "not enough rows" branch
        rows.append([float(getattr(m2, f)) for f in GEOM_FIELDS])
    out["metrics"] = np.array(rows, np.float64)
    out["pred_crc"] = np.array(crcs, np.uint32)
    # analyze_defects (:236-330) on the same maps with defect classes and pin-holes added
    from src.utils.geometry_enhanced import analyze_defects
    drows, dcrc = [], []
    for (h, w), seed, tape in GEOM_CASES:
        p = S.pred_classes(h, w, seed=seed, tape=tape, defects=True)
        dcrc.append(zlib.crc32(p.tobytes()))
        for hole_min in (10, 2):
            d = analyze_defects(p, hole_min_size=hole_min)
            drows.append([float(d.tape_hole_ratio), d.tape_num_holes, float(d.tape_coverage), d.cable_num_components,
                          d.tape_num_components, float(d.tape_largest_area_ratio), d.total_defect_area] +
                         [d.defect_areas[c] for c in (3, 4, 5, 6)])
    from src.utils.geometry_enhanced import compute_thickness_profile
    for i, ((h, w), seed, tape) in enumerate(GEOM_CASES):
        t = compute_thickness_profile(S.pred_classes(h, w, seed=seed, tape=tape, defects=True), mm_per_px=0.07)
        assert np.array_equal(t.y_coords, np.arange(h)) and t.delta_d_mm.dtype == np.float32
        out[f"profile_{i}"] = t.delta_d_mm
        out[f"profile_valid_{i}"] = np.packbits(t.valid_mask)
    # filter_cable_by_shape / constrain_tape_to_ring (src/refactor/postprocess.py) on the class masks of the same maps
    from src.refactor.postprocess import filter_cable_by_shape, constrain_tape_to_ring
    from src.refactor.config import PostprocessConfig
    for i, ((h, w), seed, tape) in enumerate(GEOM_CASES):
        p = S.pred_classes(h, w, seed=seed, tape=tape, defects=True)
        cable, tp = (p == 1).astype(np.uint8), (p == 2).astype(np.uint8)
        for name, cfg in (("default", PostprocessConfig()), ("loose", PostprocessConfig(cable_min_area=40, cable_min_aspect=1.0,
                                                                                         cable_max_center_offset=0.45, tape_ring_dilate=25, tape_ring_erode=3))):
            fc = filter_cable_by_shape(cable, cfg, w)
            out[f"cable_{name}_{i}"] = np.packbits(fc > 0)
            out[f"cable_{name}_{i}_vals"] = np.unique(fc)
            out[f"ring_{name}_{i}"] = np.packbits(constrain_tape_to_ring(tp, cable, cfg) > 0)
            out[f"ring255_{name}_{i}"] = np.packbits(constrain_tape_to_ring(tp * 255, fc, cfg) > 0)
    out["defects"] = np.array(drows, np.float64)
    out["defects_crc"] = np.array(dcrc, np.uint32)
    np.savez_compressed(os.path.join(HERE, "geometry.npz"), **out)


def viz_inputs(tag):
    """Seeded frames + masks for the visualisation vectors (regenerated identically in the tests)."""
    if tag == "full":
        fr = S.frames_cable(1, 448, 800, seed=6)[0]
        h, w = 448, 800
    else:
        fr = S.frames_noise(1, 90, 158, seed=7)[0]
        h, w = 90, 158
    roi = T.map_roi_to_original((w, h), (512, 512))
    mc = S.mask_cable_synth(h, w, roi)
    mt = (S.mask_blobs(h, w, seed=8, fill=0.02) & (1 - mc)).astype(np.uint8)
    mb = S.mask_blobs(h, w, seed=9, fill=0.004)
    return fr, mc, mt, mb, roi


def viz():
    """visualize_two_stage (infer_two_stage_burr.py:130-170) and visualize_detection (infer_enhanced_burr.py:150-192)."""
    import zlib
    out = {"versions": VERS}
    for tag in ("full", "small"):
        fr, mc, mt, mb, roi = viz_inputs(tag)
        a = T.visualize_two_stage(fr, mc, mt, mb, roi, draw_roi_box=True)
        b = T.visualize_two_stage(fr, mc, mt, mb, roi, draw_roi_box=False)
        e = E.visualize_detection(fr, mc, mt, mb, dict(zip(("x1", "y1", "x2", "y2"), roi)))
        assert np.array_equal(a, e)
        out[f"{tag}_in_crc"] = np.array(zlib.crc32(fr.tobytes() + mc.tobytes() + mt.tobytes() + mb.tobytes()), np.uint32)
        out[f"{tag}_box_crc"] = np.array(zlib.crc32(a.tobytes()), np.uint32)
        out[f"{tag}_nobox_crc"] = np.array(zlib.crc32(b.tobytes()), np.uint32)
        if tag == "small":
            out["small_box"] = a
        else:
            out["full_box_crop"] = a[160:224, 200:296].copy()
    np.savez_compressed(os.path.join(HERE, "viz.npz"), **out)


if __name__ == "__main__":
    which = sys.argv[1:] or ["stage2", "net", "enhance", "viz", "rulebased", "geometry"]
    if "stage2" in which:
        stage2()
    if "net" in which:
        net()
    if "enhance" in which:
        enhance()
    if "viz" in which:
        viz()
    if "rulebased" in which:
        rulebased()
    if "geometry" in which:
        geometry()
    for f in ("stage2.npz", "net.npz", "enhance.npz", "viz.npz", "rulebased.npz", "geometry.npz"):
        print(f, os.path.getsize(os.path.join(HERE, f)))
