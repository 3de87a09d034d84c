"""Pins oracle/ (the CPU restatement) to outputs of the REAL reference captured in
tests/golden/*.npz by tests/golden/make_golden.py."""
import cv2
import numpy as np
import pytest
import torch

import unet_b200.synth as S
from oracle import cvnp
from oracle import ref_pipeline as R

TAGS = ["cable", "noise", "speckle", "dots"]


@pytest.mark.parametrize("tag", TAGS)
def test_burr_functions_match_reference(tag, golden_stage2, golden_frames):
    g = golden_stage2
    fr = golden_frames[tag]
    gray = cv2.cvtColor(fr, cv2.COLOR_BGR2GRAY)
    mask = g[f"{tag}_mask"]
    for sens in ("medium", "high"):
        got = R.burrs_standard(gray, mask, R.BURR_CONFIGS[sens])
        assert np.array_equal(got, g[f"{tag}_burr_{sens}"])
    assert np.array_equal(R.burrs_enhanced(gray, mask), g[f"{tag}_burr_enh"])
    assert np.array_equal(R.burrs_standard(gray, np.zeros_like(mask)), g[f"{tag}_burr_empty"])
    h, w = fr.shape[:2]
    assert tuple(g[f"{tag}_roi"]) == R.map_roi((w, h))


@pytest.mark.parametrize("tag", ["cable", "noise"])
def test_preprocess_and_gray_match_reference(tag, golden_stage2, golden_frames):
    g = golden_stage2
    fr = golden_frames[tag]
    want = g[f"{tag}_pre_u8"].astype(np.float32) / 255.0
    assert np.array_equal(R.preprocess(fr).numpy(), want)
    assert np.array_equal(cvnp.bgr2gray(fr), g[f"{tag}_gray"])
    rs = cvnp.resize_linear(fr[..., ::-1], (512, 512))
    assert np.array_equal(np.transpose(rs, (2, 0, 1)), g[f"{tag}_pre_u8"])


def test_enhance_and_normalize_match_reference(golden_stage2):
    g = golden_stage2
    assert np.array_equal(R.enhance(g["enh_in"]), g["enh_out"])
    assert np.array_equal(R.normalize_frame(g["norm_in"], rotate=True), g["norm_out"])
    rot = np.ascontiguousarray(np.transpose(g["norm_in"], (1, 0, 2))[::-1])
    assert np.array_equal(cvnp.resize_linear(rot, (800, 448)), g["norm_out"])


def test_state_dict_layout(golden_net):
    sd = S.make_state_dict(0)
    assert list(sd.keys()) == list(golden_net["state_dict_keys"])
    assert len(sd) == 134


@pytest.mark.parametrize("tag", ["a", "b"])
def test_unet_forward_matches_reference(tag, golden_net):
    sd = S.make_state_dict(0)
    y = R.unet_forward(torch.from_numpy(golden_net[f"x_{tag}"]), sd).numpy()
    want = golden_net[f"y_{tag}"]
    assert y.shape == want.shape
    assert np.abs(y - want).max() <= 1e-5 * np.abs(want).max()


def test_frame_step_matches_reference(golden_net):
    torch.set_num_threads(4)
    g = golden_net
    sd = S.make_state_dict(0)
    fr = S.frames_cable(1, 448, 800, seed=int(g["step_seed"]))[0]
    st = {}
    res = R.frame_step(fr, sd, "medium", stages=st)
    assert np.abs(st["logits"][:, ::8, ::8] - g["step_logits_sub"]).max() <= 1e-4 * np.abs(g["step_logits_sub"]).max()
    agree = (res["pred"] == g["step_pred"]).mean()
    assert agree >= 0.9999          # thread-count dependent fp32 summation order in MKLDNN
    if agree == 1.0:
        assert np.array_equal(res["mask_cable"], g["step_mask_cable"])
        assert np.array_equal(res["mask_tape"], g["step_mask_tape"])
        assert np.array_equal(res["mask_burr"], g["step_mask_burr"])
        assert [res["cable_px"], res["tape_px"], res["burr_px"]] == list(g["step_counts"])


@pytest.mark.parametrize("tag", ["full", "small"])
def test_visualisation_matches_reference(golden_viz, tag):
    """oracle visualize_two_stage == the real reference's visualize_two_stage / visualize_detection."""
    import zlib
    from conftest import viz_inputs
    fr, mc, mt, mb, roi = viz_inputs(tag)
    g = golden_viz
    assert zlib.crc32(fr.tobytes() + mc.tobytes() + mt.tobytes() + mb.tobytes()) == int(g[f"{tag}_in_crc"])
    box = R.visualize_two_stage(fr, mc, mt, mb, roi, True)
    nobox = R.visualize_two_stage(fr, mc, mt, mb, roi, False)
    assert zlib.crc32(box.tobytes()) == int(g[f"{tag}_box_crc"])
    assert zlib.crc32(nobox.tobytes()) == int(g[f"{tag}_nobox_crc"])
    if tag == "small":
        assert np.array_equal(box, g["small_box"])
    else:
        assert np.array_equal(box[160:224, 200:296], g["full_box_crop"])


def test_rulebased_detector_matches_reference(golden_rulebased):
    """oracle burrs_rulebased == the real get_burr_mask_rulebased (src/refactor/burr_detector.py:11-66)."""
    from conftest import rulebased_cases, RULEBASED_CFGS
    for tag, gray, mask in rulebased_cases():
        for name, cfg in RULEBASED_CFGS.items():
            out = R.burrs_rulebased(gray, mask, **cfg)
            assert set(np.unique(out)) <= {0, 255}
            assert np.array_equal(np.packbits(out > 0), golden_rulebased[f"{tag}_{name}"]), (tag, name)
        assert R.burrs_rulebased(gray, np.zeros_like(mask)).max() == 0


@pytest.mark.parametrize("use_cv2", [True, False])
def test_diameter_metrics_match_reference(golden_geometry, use_cv2):
    """oracle diameter_metrics (cv2 filter / numpy restatement of it) == the real compute_diameter_metrics, every field."""
    from conftest import geometry_cases, GEOM_FIELDS
    for pred, want, want_short in geometry_cases(golden_geometry):
        got = R.diameter_metrics(pred, use_cv2=use_cv2)
        assert [float(got[f]) for f in GEOM_FIELDS] == list(want), (pred.shape, got)
        got = R.diameter_metrics(pred, mm_per_px=0.1, min_valid_rows=1000, use_cv2=use_cv2)
        assert [float(got[f]) for f in GEOM_FIELDS] == list(want_short)


def test_analyze_defects_matches_reference(golden_geometry):
    """oracle analyze_defects == the real analyze_defects (src/utils/geometry_enhanced.py:236-330), every field."""
    from conftest import defect_cases, defect_row
    for pred, hole_min, want in defect_cases(golden_geometry):
        assert defect_row(R.analyze_defects(pred, hole_min_size=hole_min)) == list(want), (pred.shape, hole_min)


@pytest.mark.parametrize("use_cv2", [True, False])
def test_thickness_profile_matches_reference(golden_geometry, use_cv2):
    """oracle thickness_profile == the real compute_thickness_profile (src/utils/geometry_enhanced.py:190-228)."""
    import unet_b200.synth as S
    g = golden_geometry
    for i, (h, w, seed, tape) in enumerate(g["cases"]):
        pred = S.pred_classes(int(h), int(w), seed=int(seed), tape=bool(tape), defects=True)
        d, valid = R.thickness_profile(pred, mm_per_px=0.07, use_cv2=use_cv2)
        assert d.dtype == np.float32 and np.array_equal(d, g[f"profile_{i}"]), i
        assert np.array_equal(np.packbits(valid), g[f"profile_valid_{i}"]), i


def test_postprocess_filters_match_reference(golden_geometry):
    """oracle filter_cable_by_shape / constrain_tape_to_ring == the real functions (src/refactor/postprocess.py:12-118)."""
    from conftest import postprocess_cases
    g = golden_geometry
    for i, name, cfg, cable, tape, w in postprocess_cases(g):
        fc = R.filter_cable_by_shape(cable, cfg["cable_min_area"], cfg["cable_min_aspect"], cfg["cable_max_center_offset"], w)
        assert np.array_equal(np.packbits(fc > 0), g[f"cable_{name}_{i}"]) and np.array_equal(np.unique(fc), g[f"cable_{name}_{i}_vals"])
        assert np.array_equal(np.packbits(R.constrain_tape_to_ring(tape, cable, cfg["tape_ring_dilate"], cfg["tape_ring_erode"]) > 0), g[f"ring_{name}_{i}"])
        assert np.array_equal(np.packbits(R.constrain_tape_to_ring(tape * 255, fc, cfg["tape_ring_dilate"], cfg["tape_ring_erode"]) > 0), g[f"ring255_{name}_{i}"])
