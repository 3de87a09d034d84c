import os
import sys
import zlib

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden_stage2():
    return np.load(os.path.join(GOLDEN, "stage2.npz"))


@pytest.fixture(scope="session")
def golden_net():
    return np.load(os.path.join(GOLDEN, "net.npz"))


@pytest.fixture(scope="session")
def golden_frames(golden_stage2):
    """Frames behind tests/golden/stage2.npz, regenerated from unet_b200.synth and pinned
    by the crc32 recorded when the real reference produced the fixture."""
    import unet_b200.synth as S
    g = golden_stage2
    speckle = S.frames_speckle(1, 448, 800, seed=5, n_blobs=400)[0]
    frames = {"cable": S.frames_cable(1, 448, 800, seed=0)[0], "noise": g["noise_frame"],
              "speckle": speckle, "dots": speckle}
    for k, fr in frames.items():
        assert zlib.crc32(fr.tobytes()) == int(g[f"{k}_crc"]), \
            f"synthetic frame '{k}' no longer matches the golden fixture; rerun tests/golden/make_golden.py"
    return frames


@pytest.fixture(scope="session")
def golden_viz():
    return np.load(os.path.join(GOLDEN, "viz.npz"))


def viz_inputs(tag):
    """Inputs behind tests/golden/viz.npz (same construction as make_golden.viz_inputs; the crc pins them)."""
    import unet_b200.synth as S
    from oracle import ref_pipeline as R
    if tag == "full":
        fr, (h, w) = S.frames_cable(1, 448, 800, seed=6)[0], (448, 800)
    else:
        fr, (h, w) = S.frames_noise(1, 90, 158, seed=7)[0], (90, 158)
    roi = R.map_roi((w, h))
    mc = S.mask_cable_synth(h, w, roi)
    mt = (S.mask_blobs(h, w, seed=8, fill=0.02) & (1 - mc)).astype(np.uint8)
    mb = S.mask_blobs(h, w, seed=9, fill=0.004)
    return fr, mc, mt, mb, roi


@pytest.fixture(scope="session")
def golden_rulebased():
    return np.load(os.path.join(GOLDEN, "rulebased.npz"))


def rulebased_cases():
    """(tag, gray, mask) behind tests/golden/rulebased.npz."""
    import cv2
    import unet_b200.synth as S
    for tag, fr, mask in (("speckle", S.frames_speckle(1, 448, 800, seed=5, n_blobs=400)[0], S.mask_cable_synth(448, 800, None)),
                          ("noise", S.frames_noise(1, 120, 184, seed=1)[0], S.mask_cable_synth(120, 184, None))):
        yield tag, cv2.cvtColor(fr, cv2.COLOR_BGR2GRAY), mask


RULEBASED_CFGS = {"default": dict(band_out=10, laplacian_threshold=30, min_area=20, max_area=500),
                  "wide": dict(band_out=12, laplacian_threshold=20, min_area=5, max_area=2000)}


GEOM_FIELDS = ("dc_px", "dt_px", "delta_d_px", "dc_mm", "dt_mm", "delta_d_mm", "valid_rows", "cable_coverage", "tape_coverage")


@pytest.fixture(scope="session")
def golden_geometry():
    return np.load(os.path.join(GOLDEN, "geometry.npz"))


def geometry_cases(g):
    """(pred, golden row with defaults, golden row of the not-enough-rows call) behind tests/golden/geometry.npz."""
    import unet_b200.synth as S
    for i, (h, w, seed, tape) in enumerate(g["cases"]):
        p = S.pred_classes(int(h), int(w), seed=int(seed), tape=bool(tape))
        assert zlib.crc32(p.tobytes()) == int(g["pred_crc"][i]), "synth.pred_classes changed; rerun tests/golden/make_golden.py geometry"
        yield p, g["metrics"][2 * i], g["metrics"][2 * i + 1]


def defect_cases(g):
    """(pred, hole_min_size, golden row) behind tests/golden/geometry.npz['defects']."""
    import unet_b200.synth as S
    for i, (h, w, seed, tape) in enumerate(g["cases"]):
        p = S.pred_classes(int(h), int(w), seed=int(seed), tape=bool(tape), defects=True)
        assert zlib.crc32(p.tobytes()) == int(g["defects_crc"][i])
        for j, hole_min in enumerate((10, 2)):
            yield p, hole_min, g["defects"][2 * i + j]


def defect_row(d):
    get = (lambda k: d[k]) if isinstance(d, dict) else (lambda k: getattr(d, k))
    return [float(get("tape_hole_ratio")), get("tape_num_holes"), float(get("tape_coverage")), get("cable_num_components"),
            get("tape_num_components"), float(get("tape_largest_area_ratio")), get("total_defect_area")] + [get("defect_areas")[c] for c in (3, 4, 5, 6)]


POST_CFGS = {"default": dict(cable_min_area=1000, cable_min_aspect=1.6, cable_max_center_offset=0.3, tape_ring_dilate=15, tape_ring_erode=5),
             "loose": dict(cable_min_area=40, cable_min_aspect=1.0, cable_max_center_offset=0.45, tape_ring_dilate=25, tape_ring_erode=3)}


def postprocess_cases(g):
    """(i, name, cfg, cable u8, tape u8, width) behind the cable_* / ring_* entries of tests/golden/geometry.npz."""
    import unet_b200.synth as S
    for i, (h, w, seed, tape) in enumerate(g["cases"]):
        p = S.pred_classes(int(h), int(w), seed=int(seed), tape=bool(tape), defects=True)
        for name, cfg in POST_CFGS.items():
            yield i, name, cfg, (p == 1).astype(np.uint8), (p == 2).astype(np.uint8), int(w)
