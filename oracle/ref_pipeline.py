"""ORACLE (test infrastructure, not product code).

CPU restatement of the reference's per-frame inspection path, calling the same
third-party libraries the reference calls (torch CPU, OpenCV, numpy -- all un-vendored
and unpinned in /root/reference/requirements.txt:1-4; versions installed in this
image: torch 2.11.0, opencv 4.13.0, numpy 2.3.5).  Every function cites the reference
lines it follows.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this package; the product (unet_b200) never does.

Parity pin: checked bit-exactly against outputs of the real reference imported from
/root/reference (tests/golden/make_golden.py -> tests/golden/*.npz,
tests/test_oracle_golden.py).  The reference itself ships no tests or golden vectors.
"""
from __future__ import annotations

import cv2
import numpy as np
import torch
import torch.nn.functional as F

# ---------------------------------------------------------------------------
# Stage 1: the network (src/models/unetpp.py)
# ---------------------------------------------------------------------------

ENCODER = ["conv0_0", "conv1_0", "conv2_0", "conv3_0", "conv4_0"]
DECODER = ["conv3_1", "conv2_2", "conv1_3", "conv0_4"]


def _conv_bn_relu(x, sd, conv, bn):
    # ConvBlock.forward, src/models/unetpp.py:23-26 (eval-mode BN, eps 1e-5)
    y = F.conv2d(x, sd[conv + ".weight"], sd[conv + ".bias"], padding=1)
    y = F.batch_norm(y, sd[bn + ".running_mean"], sd[bn + ".running_var"],
                     sd[bn + ".weight"], sd[bn + ".bias"], training=False, eps=1e-5)
    return F.relu(y)


def _block(x, sd, name):
    x = _conv_bn_relu(x, sd, name + ".conv1", name + ".bn1")
    return _conv_bn_relu(x, sd, name + ".conv2", name + ".bn2")


def unet_forward(x: torch.Tensor, sd: dict, return_features: bool = False):
    """Eval-mode NestedUNet.forward (src/models/unetpp.py:93-135) as a function of the
    state_dict: 5 encoder blocks with 2x2 max-pool (:104-108), 4 diagonal decoder
    blocks over cat([skip, bilinear-x2(align_corners=True)(below)]) (:111-116),
    1x1 head (:119).  fp32, CPU."""
    feats = {}
    with torch.no_grad():
        skips = []
        h = x
        for i, name in enumerate(ENCODER):
            if i:
                h = F.max_pool2d(h, 2, 2)
            h = _block(h, sd, name)
            feats[name] = h
            skips.append(h)
        for name, skip in zip(DECODER, reversed(skips[:-1])):
            up = F.interpolate(h, scale_factor=2, mode="bilinear", align_corners=True)
            h = _block(torch.cat([skip, up], dim=1), sd, name)
            feats[name] = h
        out = F.conv2d(h, sd["final.weight"], sd["final.bias"])
    return (out, feats) if return_features else out


# ---------------------------------------------------------------------------
# Pre / post processing around the net (infer_two_stage_burr.py)
# ---------------------------------------------------------------------------

FIXED_ROI_512 = (140, 0, 270, 512)        # infer_two_stage_burr.py:29-34 (x1,y1,x2,y2)
VERTICAL_ROI = (200, 0, 600, 448)         # infer_enhanced_burr.py:23-28

BURR_CONFIGS = {                          # infer_two_stage_burr.py:194-198 (only
    "low": {"min_area": 50, "max_area": 800},       # min/max_area are ever read, :114)
    "medium": {"min_area": 30, "max_area": 800},
    "high": {"min_area": 20, "max_area": 1000},
}
ENHANCED_CONFIG = {"min_area": 50, "max_area": 500}   # infer_enhanced_burr.py:208-212


def map_roi(orig_wh, target_wh=(512, 512)):
    """map_roi_to_original, infer_two_stage_burr.py:37-47."""
    sx = orig_wh[0] / target_wh[0]
    sy = orig_wh[1] / target_wh[1]
    x1, y1, x2, y2 = FIXED_ROI_512
    return int(x1 * sx), int(y1 * sy), int(x2 * sx), int(y2 * sy)


def normalize_frame(frame, rotate=False, size=(800, 448)):
    """infer_two_stage_burr.py:275-280: optional 90-degree CCW rotate, LINEAR resize."""
    if rotate:
        frame = cv2.rotate(frame, cv2.ROTATE_90_COUNTERCLOCKWISE)
    return cv2.resize(frame, size, interpolation=cv2.INTER_LINEAR)


def preprocess(frame_bgr: np.ndarray, size=(512, 512)) -> torch.Tensor:
    """preprocess_image, infer_two_stage_burr.py:122-127."""
    rgb = cv2.cvtColor(frame_bgr, cv2.COLOR_BGR2RGB)
    small = cv2.resize(rgb, size, interpolation=cv2.INTER_LINEAR)
    chw = np.transpose(small.astype(np.float32) / 255.0, (2, 0, 1))
    return torch.from_numpy(np.ascontiguousarray(chw)).float()


def masks_from_logits(logits: torch.Tensor, frame_wh, roi):
    """infer_two_stage_burr.py:299-314: softmax -> argmax -> class masks -> NEAREST
    upscale -> keep only the ROI.  logits [3,h,w] fp32."""
    probs = torch.softmax(logits[None], dim=1)[0].cpu().numpy()
    pred = np.argmax(probs, axis=0).astype(np.uint8)
    out = []
    x1, y1, x2, y2 = roi
    for cls in (1, 2):
        m = (pred == cls).astype(np.uint8)
        full = cv2.resize(m, frame_wh, interpolation=cv2.INTER_NEAREST)
        keep = np.zeros_like(full)
        keep[y1:y2, x1:x2] = full[y1:y2, x1:x2]
        out.append(keep)
    return pred, out[0], out[1]


# ---------------------------------------------------------------------------
# Stage 2: burr detection
# ---------------------------------------------------------------------------

def _ellipse(k):
    return cv2.getStructuringElement(cv2.MORPH_ELLIPSE, (k, k))


def _band(mask, k):
    dil = cv2.dilate(mask, _ellipse(k), iterations=1)
    return dil & (~mask)


def _filter_components(cand, shape_like, min_area, max_area, max_aspect, min_side,
                       stages=None):
    n, labels, stats, _ = cv2.connectedComponentsWithStats(cand, connectivity=8)
    out = np.zeros_like(shape_like)
    for i in range(1, n):
        area = stats[i, cv2.CC_STAT_AREA]
        w = stats[i, cv2.CC_STAT_WIDTH]
        h = stats[i, cv2.CC_STAT_HEIGHT]
        aspect = max(w, h) / (min(w, h) + 1e-6)
        if min_area <= area <= max_area and aspect < max_aspect and w >= min_side and h >= min_side:
            out[labels == i] = 1
    if stages is not None:
        stages["labels"] = labels
        stages["stats"] = stats
    return out


def burrs_standard(gray, mask_cable, config=None, stages=None):
    """detect_burrs_on_cable, infer_two_stage_burr.py:50-119.  `stages`, when a dict,
    receives every intermediate for stage-by-stage kernel parity."""
    cfg = config or {"min_area": 30, "max_area": 800}          # :61-67
    if mask_cable.max() == 0:                                    # :69-70
        return np.zeros_like(mask_cable)
    contours, _ = cv2.findContours(mask_cable, cv2.RETR_EXTERNAL, cv2.CHAIN_APPROX_SIMPLE)   # :73-75 (never empty here; kept so
    if len(contours) == 0:                                       # that the timed reference arm does the reference's work)
        return np.zeros_like(mask_cable)
    band = _band(mask_cable, 8)                                  # :78-81
    blur = cv2.GaussianBlur(gray, (5, 5), 1.0)                   # :85
    edges = cv2.Canny(blur, 50, 150)                             # :86
    cand = edges & band                                          # :89
    closed = cv2.morphologyEx(cand, cv2.MORPH_CLOSE, _ellipse(3))    # :92-93
    opened = cv2.morphologyEx(closed, cv2.MORPH_OPEN, _ellipse(2))   # :96-97
    if stages is not None:
        stages.update(band=band, blur=blur, edges=edges, cand=cand, closed=closed,
                      opened=opened)
    # :100-117  (w > 3 and h > 3  <=>  min side 4)
    return _filter_components(opened, mask_cable, cfg["min_area"], cfg["max_area"],
                              5.0, 4, stages)


def burrs_enhanced(gray, mask_cable, config=None, stages=None):
    """detect_burrs_enhanced, infer_enhanced_burr.py:69-138."""
    cfg = config or ENHANCED_CONFIG
    if mask_cable.max() == 0:                                    # :73-74
        return np.zeros_like(mask_cable)
    contours, _ = cv2.findContours(mask_cable, cv2.RETR_EXTERNAL, cv2.CHAIN_APPROX_SIMPLE)   # :77-79
    if len(contours) == 0:
        return np.zeros_like(mask_cable)
    band = _band(mask_cable, 25)                                 # :82-85
    blur = cv2.GaussianBlur(gray, (5, 5), 1.0)                   # :89
    e_canny = cv2.Canny(blur, 30, 100)                           # :90
    sx = cv2.Sobel(gray, cv2.CV_64F, 1, 0, ksize=3)              # :93-94
    sy = cv2.Sobel(gray, cv2.CV_64F, 0, 1, ksize=3)
    mag = np.sqrt(sx ** 2 + sy ** 2)                             # :95
    with np.errstate(invalid="ignore", divide="ignore"):
        mag8 = np.uint8(mag / mag.max() * 255)                   # :96 (truncating cast)
    e_sobel = cv2.threshold(mag8, 50, 255, cv2.THRESH_BINARY)[1]     # :97
    lap = cv2.Laplacian(gray, cv2.CV_64F)                        # :100
    lap8 = np.abs(lap).astype(np.uint8)                          # :101 (wraps mod 256)
    e_lap = cv2.threshold(lap8, 15, 255, cv2.THRESH_BINARY)[1]   # :102
    fused = cv2.bitwise_or(cv2.bitwise_or(e_canny, e_sobel), e_lap)  # :105-106
    cand = fused & band                                          # :109
    closed = cv2.morphologyEx(cand, cv2.MORPH_CLOSE, _ellipse(5))    # :112-113
    opened = cv2.morphologyEx(closed, cv2.MORPH_OPEN, _ellipse(3))   # :115-116
    if stages is not None:
        stages.update(band=band, blur=blur, edges_canny=e_canny, edges_sobel=e_sobel,
                      edges_lap=e_lap, cand=cand, closed=closed, opened=opened)
    # :119-136  (aspect < 6, w >= 5, h >= 5)
    return _filter_components(opened, mask_cable, cfg["min_area"], cfg["max_area"],
                              6.0, 5, stages)


def enhance(frame_bgr, stages=None):
    """enhance_image, infer_enhanced_burr.py:38-66."""
    lab = cv2.cvtColor(frame_bgr, cv2.COLOR_BGR2LAB)             # :46
    l, a, b = cv2.split(lab)
    l2 = cv2.createCLAHE(clipLimit=3.0, tileGridSize=(8, 8)).apply(l)   # :50-51
    bgr = cv2.cvtColor(cv2.merge([l2, a, b]), cv2.COLOR_LAB2BGR)        # :54-55
    den = cv2.fastNlMeansDenoisingColored(bgr, None, 10, 10, 7, 21)     # :58
    k = np.array([[-1, -1, -1], [-1, 9, -1], [-1, -1, -1]])             # :61-63
    out = cv2.filter2D(den, -1, k)                                      # :64
    if stages is not None:
        stages.update(lab=lab, l_clahe=l2, bgr_clahe=bgr, denoised=den)
    return out


# ---------------------------------------------------------------------------
# Whole per-frame step (the body of the while-loop, infer_two_stage_burr.py:286-343)
# ---------------------------------------------------------------------------

def frame_step(frame_bgr, sd, sensitivity="medium", variant="standard", roi=None,
               do_enhance=False, stages=None):
    """One iteration of the reference's hot loop without video I/O / drawing.
    Returns dict(cable_px, tape_px, burr_px, mask_cable, mask_tape, mask_burr, pred)."""
    h, w = frame_bgr.shape[:2]
    if variant == "enhanced":
        if do_enhance:
            frame_bgr = enhance(frame_bgr)                      # infer_enhanced_burr.py:279-280
        roi = roi or VERTICAL_ROI                               # :304
    else:
        roi = roi or map_roi((w, h))                            # infer_two_stage_burr.py:253
    x = preprocess(frame_bgr)[None]                             # :292
    logits = unet_forward(x, sd)[0]                             # :294-297
    pred, m_cable, m_tape = masks_from_logits(logits, (w, h), roi)   # :299-314
    gray = cv2.cvtColor(frame_bgr, cv2.COLOR_BGR2GRAY)          # :317
    if variant == "enhanced":
        m_burr = burrs_enhanced(gray, m_cable, ENHANCED_CONFIG, stages)
    else:
        m_burr = burrs_standard(gray, m_cable, BURR_CONFIGS[sensitivity], stages)   # :318
    if stages is not None:
        stages.update(gray=gray, logits=logits.numpy())
    return dict(cable_px=int(m_cable.sum()), tape_px=int(m_tape.sum()),
                burr_px=int(m_burr.sum()), mask_cable=m_cable, mask_tape=m_tape,
                mask_burr=m_burr, pred=pred, roi=roi)


def stats_line(idx, frame_no, total, cable_px, tape_px, burr_px, roi):
    """Progress line, infer_two_stage_burr.py:333-343,358-359."""
    x1, y1, x2, y2 = roi
    area = (x2 - x1) * (y2 - y1)
    r = [(p / area * 100 if area > 0 else 0) for p in (cable_px, tape_px, burr_px)]
    status = "[BURR!]" if burr_px > 0 else "[OK]"
    return (f"[{idx:4d}] Frame {frame_no}/{total} {status} | "
            f"Cable:{r[0]:5.1f}% Tape:{r[1]:5.1f}% Burr:{r[2]:5.1f}%")


def visualize_two_stage(frame, mask_cable, mask_tape, mask_burr, roi, draw_roi_box=True):
    """visualize_two_stage (infer_two_stage_burr.py:130-170) = visualize_detection (infer_enhanced_burr.py:150-192):
    the addWeighted chain restated in oracle/cvnp.py (overlay_blend), then cv2's own drawing calls."""
    from . import cvnp
    x1, y1, x2, y2 = roi
    out = np.ascontiguousarray(cvnp.overlay_blend(frame, mask_cable, mask_tape, mask_burr, roi))
    if draw_roi_box:
        cv2.rectangle(out, (x1, y1), (x2, y2), (0, 255, 255), 2)
        cv2.putText(out, "ROI", (x1 + 5, y1 + 25), cv2.FONT_HERSHEY_SIMPLEX, 0.6, (0, 255, 255), 2)
    for mask, colour, thick in ((mask_cable, (0, 255, 0), 2), (mask_tape, (255, 0, 0), 2), (mask_burr, (255, 0, 255), 3)):
        contours, _ = cv2.findContours(np.ascontiguousarray(mask), cv2.RETR_EXTERNAL, cv2.CHAIN_APPROX_SIMPLE)
        cv2.drawContours(out, contours, -1, colour, thick)
    return out


def burrs_rulebased(gray, mask_cable, band_out=10, laplacian_threshold=30, min_area=20, max_area=500):
    """get_burr_mask_rulebased, src/refactor/burr_detector.py:11-66 (BurrConfig defaults, src/refactor/config.py:98-103)."""
    if mask_cable.max() == 0:                                                        # :23-24
        return np.zeros_like(mask_cable)
    k = 2 * band_out + 1
    dil = cv2.dilate(mask_cable, _ellipse(k), iterations=1)                          # :37-38
    band = cv2.subtract(dil, mask_cable)                                             # :41
    lap8 = np.abs(cv2.Laplacian(gray, cv2.CV_64F)).astype(np.uint8)                  # :44-45 (wraps mod 256)
    lap_band = cv2.bitwise_and(lap8, lap8, mask=band)                                # :48
    burr = cv2.threshold(lap_band, laplacian_threshold, 255, cv2.THRESH_BINARY)[1]   # :51
    n, labels, stats, _ = cv2.connectedComponentsWithStats(burr.astype(np.uint8), connectivity=8)   # :54-56
    out = np.zeros_like(burr)
    for i in range(1, n):                                                            # :60-64
        if min_area <= stats[i, cv2.CC_STAT_AREA] <= max_area:
            out[labels == i] = 255
    return out


def _largest_cc(mask, min_area):
    """_largest_connected_component, src/utils/geometry_enhanced.py:79-110."""
    n, labels, stats, _ = cv2.connectedComponentsWithStats(mask, connectivity=8)
    if n <= 1:
        return mask
    areas = stats[1:, cv2.CC_STAT_AREA]
    ok = np.where(areas >= min_area)[0]
    if len(ok) == 0:
        return np.zeros_like(mask)
    return (labels == 1 + int(ok[np.argmax(areas[ok])])).astype(np.uint8)


def _row_widths(mask, use_cv2=True):
    """_compute_width_per_row(mask, smooth=True, kernel_size=31), src/utils/geometry_enhanced.py:45-76."""
    from . import cvnp
    h = mask.shape[0]
    w = np.zeros(h, np.float32)
    for y in range(h):
        xs = np.where(mask[y] > 0)[0]
        if xs.size:
            w[y] = float(xs.max() - xs.min() + 1)
    if use_cv2:
        return cv2.GaussianBlur(w.reshape(-1, 1), (1, 31), sigmaX=0).reshape(-1)
    return cvnp.gauss31_f32(w)


def diameter_metrics(pred, cable_cls=1, tape_cls=2, mm_per_px=0.05, min_valid_rows=20, use_cv2=True):
    """compute_diameter_metrics, src/utils/geometry_enhanced.py:113-187 -> dict with the dataclass's fields."""
    h, w = pred.shape
    cable = _largest_cc((pred == cable_cls).astype(np.uint8), 50)
    tape = _largest_cc((pred == tape_cls).astype(np.uint8), 50)
    cw, tw = _row_widths(cable, use_cv2), _row_widths(tape, use_cv2)
    valid = (cw > 0) & (tw > 0)
    out = dict(valid_rows=int(valid.sum()), cable_coverage=cable.sum() / (h * w), tape_coverage=tape.sum() / (h * w))
    if out["valid_rows"] < min_valid_rows:
        out.update(dc_px=0.0, dt_px=0.0, delta_d_px=0.0, dc_mm=0.0, dt_mm=0.0, delta_d_mm=0.0)
        return out
    dc, dt = float(np.median(cw[valid])), float(np.median(tw[valid]))
    out.update(dc_px=dc, dt_px=dt, delta_d_px=dt - dc, dc_mm=dc * mm_per_px, dt_mm=dt * mm_per_px,
               delta_d_mm=dt * mm_per_px - dc * mm_per_px)
    return out


def analyze_defects(pred, cable_cls=1, tape_cls=2, defect_classes=(3, 4, 5, 6), hole_min_size=10):
    """analyze_defects, src/utils/geometry_enhanced.py:236-330 -> dict with the dataclass's fields."""
    h, w = pred.shape
    tape = (pred == tape_cls).astype(np.uint8)
    n_tape, _, tstats, _ = cv2.connectedComponentsWithStats(tape, connectivity=8)                    # :267-269
    closed = cv2.morphologyEx(tape, cv2.MORPH_CLOSE, _ellipse(5))                                    # :273-274
    holes = ((closed.astype(np.int16) - tape.astype(np.int16)) > 0).astype(np.uint8)                 # :275-276
    n_h, _, hstats, _ = cv2.connectedComponentsWithStats(holes, connectivity=8)                      # :279-281
    ok = [int(a) for a in (hstats[1:, cv2.CC_STAT_AREA] if n_h > 1 else []) if a >= hole_min_size]   # :283-285
    n_cable = cv2.connectedComponentsWithStats((pred == cable_cls).astype(np.uint8), connectivity=8)[0]
    areas = {int(c): int((pred == c).sum()) for c in defect_classes}                                 # :308-314
    n_tape = max(0, n_tape - 1)
    return dict(tape_hole_ratio=sum(ok) / max(int(tape.sum()), 1), tape_num_holes=len(ok), tape_coverage=tape.sum() / (h * w),
                cable_num_components=max(0, n_cable - 1), tape_num_components=n_tape,
                tape_largest_area_ratio=(tstats[1:, cv2.CC_STAT_AREA].max() / tape.sum()) if n_tape > 0 else 0.0,
                defect_areas=areas, total_defect_area=sum(areas.values()))


def thickness_profile(pred, cable_cls=1, tape_cls=2, mm_per_px=0.05, use_cv2=True):
    """compute_thickness_profile, src/utils/geometry_enhanced.py:190-228 -> (delta_d_mm float32 [H], valid bool [H])."""
    cw = _row_widths((pred == cable_cls).astype(np.uint8), use_cv2)
    tw = _row_widths((pred == tape_cls).astype(np.uint8), use_cv2)
    return (tw - cw) * mm_per_px, (cw > 0) & (tw > 0)


def filter_cable_by_shape(mask, min_area=1000, min_aspect=1.6, max_center_offset=0.3, roi_width=512):
    """filter_cable_by_shape, src/refactor/postprocess.py:12-77 (PostprocessConfig defaults, src/refactor/config.py:60-70)."""
    if mask.max() == 0:
        return mask
    n, labels, stats, cent = cv2.connectedComponentsWithStats(mask.astype(np.uint8), connectivity=8)
    if n <= 1:
        return np.zeros_like(mask)
    best_score, best = -1, -1
    for i in range(1, n):
        area, w, h = stats[i, cv2.CC_STAT_AREA], stats[i, cv2.CC_STAT_WIDTH], stats[i, cv2.CC_STAT_HEIGHT]
        if area < min_area:
            continue
        aspect = max(w, h) / (min(w, h) + 1e-6)
        if aspect < min_aspect:
            continue
        off = abs(cent[i][0] - roi_width / 2.0) / roi_width
        if off > max_center_offset:
            continue
        score = area * aspect * (1.0 - off)
        if score > best_score:
            best_score, best = score, i
    return (labels == best).astype(np.uint8) * 255 if best > 0 else np.zeros_like(mask)


def constrain_tape_to_ring(tape, cable, dilate_k=15, erode_k=5):
    """constrain_tape_to_ring, src/refactor/postprocess.py:80-118."""
    if cable.max() == 0 or tape.max() == 0:
        return np.zeros_like(tape)
    ring = cv2.subtract(cv2.dilate(cable, _ellipse(dilate_k), iterations=1), cv2.erode(cable, _ellipse(erode_k), iterations=1))
    con = cv2.bitwise_and(tape, ring)
    if con.max() > 0:
        n, labels, stats, _ = cv2.connectedComponentsWithStats(con.astype(np.uint8), connectivity=8)
        if n > 1:
            con = (labels == np.argmax(stats[1:, cv2.CC_STAT_AREA]) + 1).astype(np.uint8) * 255
    return con
