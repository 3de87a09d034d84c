"""Function contract of the reference's per-frame helpers, executed by libunetb200.

Same names, argument meaning and results as the reference functions:
  map_roi_to_original     infer_two_stage_burr.py:37-47
  preprocess_image        infer_two_stage_burr.py:122-127
  detect_burrs_on_cable   infer_two_stage_burr.py:50-119
  detect_burrs_enhanced   infer_enhanced_burr.py:69-138
  enhance_image           infer_enhanced_burr.py:38-66
  normalize_frame         infer_two_stage_burr.py:275-280 (rotate + resize)
  get_burr_mask_rulebased src/refactor/burr_detector.py:11-66 (the Laplacian-band detector of the refactored library)
  compute_diameter_metrics src/utils/geometry_enhanced.py:113-187 (Dc / Dt / delta-D of the refactored library)
  compute_thickness_profile src/utils/geometry_enhanced.py:190-228 (per-row delta-D profile)
  filter_cable_by_shape, constrain_tape_to_ring   src/refactor/postprocess.py:12-118 (shape constraints on the masks)
  analyze_defects         src/utils/geometry_enhanced.py:236-330 (coverage, components, hole ratio, defect areas)
  visualize_two_stage     infer_two_stage_burr.py:130-170 (blend, ROI box + label and the three contour sets on the device)
Each accepts what the reference accepts (one numpy frame on the host) and returns the same
kind of object; it ALSO accepts a batched CUDA tensor ([B,...]) and then returns CUDA tensors
without touching the host.  All arithmetic runs in the CUDA library; there is no CPU path.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib

FIXED_ROI_512 = {"x1": 140, "y1": 0, "x2": 270, "y2": 512}
VERTICAL_ROI = {"x1": 200, "y1": 0, "x2": 600, "y2": 448}
BURR_CONFIGS = {
    "low": {"band_out": 10, "laplacian_threshold": 35, "min_area": 50, "max_area": 800, "morph_kernel": 3},
    "medium": {"band_out": 15, "laplacian_threshold": 25, "min_area": 30, "max_area": 800, "morph_kernel": 3},
    "high": {"band_out": 20, "laplacian_threshold": 20, "min_area": 20, "max_area": 1000, "morph_kernel": 5},
}

_contexts = {}          # insertion-ordered: least recently used first
_MAX_CONTEXTS = 4


def _ctx(device, batch, frame_hw, net_hw=(16, 16)) -> _lib.Context:
    """Cached scratch context for stand-alone operator calls.  Operators that never run the network get minimal network
    planes (16x16); at most _MAX_CONTEXTS contexts are kept, the least recently used one is closed."""
    dev = torch.device(device)
    key = (dev.index or 0, tuple(frame_hw), tuple(net_hw))
    c = _contexts.pop(key, None)
    if c is None or c.max_batch < batch:
        if c is not None:
            c.close()
        while len(_contexts) >= _MAX_CONTEXTS:
            _contexts.pop(next(iter(_contexts))).close()
        c = _lib.Context(dev.index or 0, max_batch=max(batch, 1), frame_hw=frame_hw, net_hw=net_hw)
    _contexts[key] = c
    return c


def _default_device():
    if not torch.cuda.is_available():
        raise _lib.UbError("no CUDA device: unet_b200 has no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def map_roi_to_original(original_size, target_size=(512, 512)):
    orig_w, orig_h = original_size
    tw, th = target_size
    sx, sy = orig_w / tw, orig_h / th
    r = FIXED_ROI_512
    return (int(r["x1"] * sx), int(r["y1"] * sy), int(r["x2"] * sx), int(r["y2"] * sy))


def _batched(a, ndim_single):
    """-> (cuda tensor with batch dim, was_numpy, was_single)"""
    if isinstance(a, np.ndarray):
        t = torch.from_numpy(np.ascontiguousarray(a)).to(_default_device())
        return (t[None] if a.ndim == ndim_single else t).contiguous(), True, a.ndim == ndim_single
    single = a.dim() == ndim_single
    return (a[None] if single else a).contiguous(), False, single


def _unbatch(t, was_numpy, was_single):
    if was_single:
        t = t[0]
    return t.cpu().numpy() if was_numpy else t


def preprocess_image(frame, target_size=(512, 512)):
    """BGR uint8 frame [H,W,3] -> float32 tensor [3,th,tw] (RGB, /255).  A numpy frame gives a
    CPU tensor like the reference; a CUDA uint8 tensor ([H,W,3] or [B,H,W,3]) stays on the GPU."""
    f, was_np, single = _batched(frame, 3)
    tw, th = target_size
    c = _ctx(f.device, f.shape[0], tuple(f.shape[1:3]), (th, tw))
    with torch.cuda.device(f.device):
        x, _ = c.preprocess(f, want_x=True, want_gray=False)
    x = x[0] if single else x
    return x.cpu() if was_np else x


def bgr2gray(frame):
    """cv2.cvtColor(frame, COLOR_BGR2GRAY) (infer_two_stage_burr.py:317)."""
    f, was_np, single = _batched(frame, 3)
    c = _ctx(f.device, f.shape[0], tuple(f.shape[1:3]))
    with torch.cuda.device(f.device):
        _, g = c.preprocess(f, want_x=False, want_gray=True)
    return _unbatch(g, was_np, single)


def normalize_frame(frame, size=(800, 448), rotate=False):
    f, was_np, single = _batched(frame, 3)
    c = _ctx(f.device, f.shape[0], (size[1], size[0]))
    with torch.cuda.device(f.device):
        out = c.normalize_frames(f, size, rotate)
    return _unbatch(out, was_np, single)


def _detect(variant, frame_gray, mask_cable, params):
    g, was_np, single = _batched(frame_gray, 2)
    m, _, _ = _batched(mask_cable, 2)
    c = _ctx(g.device, g.shape[0], tuple(g.shape[1:3]))
    with torch.cuda.device(g.device):
        out, _ = c.burr(g, m.to(g.device), variant, params)
    return _unbatch(out, was_np, single)


def detect_burrs_on_cable(frame_gray, mask_cable, config=None):
    """Standard burr detector; only min_area / max_area of `config` are read, exactly like the
    reference (infer_two_stage_burr.py:114)."""
    cfg = config or {"min_area": 30, "max_area": 800}
    return _detect(_lib.VARIANT_STANDARD, frame_gray, mask_cable,
                   _lib.standard_params(cfg["min_area"], cfg["max_area"]))


def detect_burrs_enhanced(frame_gray, mask_cable, config):
    return _detect(_lib.VARIANT_ENHANCED, frame_gray, mask_cable,
                   _lib.enhanced_params(config["min_area"], config["max_area"]))


def get_burr_mask_rulebased(frame_gray, mask_cable, cfg=None):
    """get_burr_mask_rulebased(frame_gray, mask_cable, cfg: BurrConfig) (src/refactor/burr_detector.py:11-66):
    band = dilate(mask, ELLIPSE(2*band_out+1)) - mask, wrapped |Laplacian| > laplacian_threshold inside the band,
    components with min_area <= area <= max_area -> {0, 255}.  `cfg` is the reference's BurrConfig, any object with
    the same four attributes, a dict, or None (BurrConfig defaults 10 / 30 / 20 / 500).  band_out <= 12."""
    get = (lambda k, d: cfg.get(k, d)) if isinstance(cfg, dict) else (lambda k, d: getattr(cfg, k, d))
    p = _lib.rulebased_params(get("band_out", 10), get("laplacian_threshold", 30), get("min_area", 20), get("max_area", 500))
    if isinstance(frame_gray, np.ndarray) and frame_gray.ndim == 3 and frame_gray.shape[2] == 3:
        frame_gray = bgr2gray(frame_gray)                     # :26-27
    out = _detect(_lib.VARIANT_LAPLACIAN, frame_gray, mask_cable, p)
    return out * 255


class DiameterMetrics:
    """Same fields as the reference's dataclass (src/utils/geometry_enhanced.py:18-34)."""
    __slots__ = ("dc_px", "dt_px", "delta_d_px", "dc_mm", "dt_mm", "delta_d_mm", "valid_rows", "cable_coverage", "tape_coverage")

    def __init__(self, **kw):
        for k in self.__slots__:
            setattr(self, k, kw[k])

    def __repr__(self):
        return "DiameterMetrics(" + ", ".join(f"{k}={getattr(self, k)!r}" for k in self.__slots__) + ")"


def compute_diameter_metrics(pred_mask, cable_cls=1, tape_cls=2, mm_per_px=0.05, min_valid_rows=20):
    """compute_diameter_metrics(pred_mask, cable_cls, tape_cls, mm_per_px, min_valid_rows)
    (src/utils/geometry_enhanced.py:113-187).  pred_mask: class ids u8 [H,W] (numpy -> one DiameterMetrics) or a
    CUDA tensor [B,H,W] (-> list).  Labelling, row extents, float32 smoothing and medians run on the device; the
    scalar conversions to millimetres are the reference's own float arithmetic on the host."""
    f, was_np, single = _batched(pred_mask, 2)
    b, h, w = f.shape
    c = _ctx(f.device, b, (h, w), (16, 16))
    with torch.cuda.device(f.device):
        raw = c.diameter_metrics(f, cable_cls, tape_cls, 50, min_valid_rows).cpu().numpy()
    res = []
    for r in raw:
        dc_px, dt_px = (float(v) for v in r[:2].view(np.float32))
        valid_rows = int(r[2])
        if valid_rows < min_valid_rows:
            dc_px = dt_px = 0.0
        dc_mm, dt_mm = dc_px * mm_per_px, dt_px * mm_per_px
        res.append(DiameterMetrics(dc_px=dc_px, dt_px=dt_px, delta_d_px=dt_px - dc_px, dc_mm=dc_mm, dt_mm=dt_mm,
                                   delta_d_mm=dt_mm - dc_mm, valid_rows=valid_rows, cable_coverage=int(r[3]) / (h * w),
                                   tape_coverage=int(r[4]) / (h * w)))
    return res[0] if single else res


def _cfg_get(cfg, name, default):
    if cfg is None:
        return default
    return cfg.get(name, default) if isinstance(cfg, dict) else getattr(cfg, name, default)


def filter_cable_by_shape(mask_cable, cfg=None, roi_width=512):
    """filter_cable_by_shape(mask_cable, cfg: PostprocessConfig, roi_width) (src/refactor/postprocess.py:12-77): keep the
    component with the best area * aspect * (1 - centre offset) among those passing the three gates -> {0, 255}.
    cfg: the reference's PostprocessConfig, a dict, or None (defaults 1000 / 1.6 / 0.3, src/refactor/config.py:60-70)."""
    m, was_np, single = _batched(mask_cable, 2)
    c = _ctx(m.device, m.shape[0], tuple(m.shape[1:3]), (16, 16))
    with torch.cuda.device(m.device):
        out = c.filter_cable_by_shape(m, _cfg_get(cfg, "cable_min_area", 1000), _cfg_get(cfg, "cable_min_aspect", 1.6),
                                      _cfg_get(cfg, "cable_max_center_offset", 0.3), roi_width)
    return _unbatch(out, was_np, single)


def constrain_tape_to_ring(mask_tape, mask_cable, cfg=None):
    """constrain_tape_to_ring(mask_tape, mask_cable, cfg) (src/refactor/postprocess.py:80-118): tape restricted to
    dilate(cable, 15) - erode(cable, 5), largest component -> {0, 255}.  Binary masks ({0,1} or {0,255})."""
    t, was_np, single = _batched(mask_tape, 2)
    cb, _, _ = _batched(mask_cable, 2)
    c = _ctx(t.device, t.shape[0], tuple(t.shape[1:3]), (16, 16))
    with torch.cuda.device(t.device):
        out = c.constrain_tape_to_ring(t, cb.to(t.device), _cfg_get(cfg, "tape_ring_dilate", 15), _cfg_get(cfg, "tape_ring_erode", 5))
    return _unbatch(out, was_np, single)


class ThicknessProfile:
    """Same fields as the reference's dataclass (src/utils/geometry_enhanced.py:37-42)."""
    __slots__ = ("y_coords", "delta_d_mm", "valid_mask")

    def __init__(self, y_coords, delta_d_mm, valid_mask):
        self.y_coords, self.delta_d_mm, self.valid_mask = y_coords, delta_d_mm, valid_mask


def compute_thickness_profile(pred_mask, cable_cls=1, tape_cls=2, mm_per_px=0.05):
    """compute_thickness_profile(pred_mask, cable_cls, tape_cls, mm_per_px) (src/utils/geometry_enhanced.py:190-228).
    numpy [H,W] -> one ThicknessProfile of numpy arrays; CUDA tensor [B,H,W] -> list.  Row extents and the float32
    smoothing run on the device; the subtraction and the scale are the reference's own numpy expressions."""
    f, was_np, single = _batched(pred_mask, 2)
    b, h, w = f.shape
    c = _ctx(f.device, b, (h, w), (16, 16))
    with torch.cuda.device(f.device):
        prof = c.width_profiles(f, cable_cls, tape_cls).cpu().numpy()
    res = []
    for p in prof:
        cw, tw = p[0], p[1]
        delta_d_px = tw - cw                                          # :215
        res.append(ThicknessProfile(np.arange(h), delta_d_px * mm_per_px, (cw > 0) & (tw > 0)))   # :216-226
    return res[0] if single else res


class DefectAnalysis:
    """Same fields as the reference's dataclass (src/utils/geometry_enhanced.py:216-233)."""
    __slots__ = ("tape_hole_ratio", "tape_num_holes", "tape_coverage", "cable_num_components", "tape_num_components",
                 "tape_largest_area_ratio", "defect_areas", "total_defect_area")

    def __init__(self, **kw):
        for k in self.__slots__:
            setattr(self, k, kw[k])

    def __repr__(self):
        return "DefectAnalysis(" + ", ".join(f"{k}={getattr(self, k)!r}" for k in self.__slots__) + ")"


def analyze_defects(pred_mask, cable_cls=1, tape_cls=2, defect_classes=(3, 4, 5, 6), hole_min_size=10):
    """analyze_defects(pred_mask, cable_cls, tape_cls, defect_classes, hole_min_size)
    (src/utils/geometry_enhanced.py:236-330).  numpy [H,W] -> one DefectAnalysis; CUDA tensor [B,H,W] -> list.
    Class ids must be < 8.  Labelling, the 5x5 CLOSE, hole components and the class histogram run on the device."""
    f, was_np, single = _batched(pred_mask, 2)
    b, h, w = f.shape
    c = _ctx(f.device, b, (h, w), (16, 16))
    with torch.cuda.device(f.device):
        raw = c.analyze_defects(f, cable_cls, tape_cls, hole_min_size).cpu().numpy()
    res = []
    for r in raw:
        tape_px, tape_n, tape_largest, holes, hole_px, cable_n = (int(v) for v in r[:6])
        areas = {int(cls): int(r[8 + cls]) for cls in defect_classes}
        res.append(DefectAnalysis(tape_hole_ratio=hole_px / max(tape_px, 1), tape_num_holes=holes, tape_coverage=tape_px / (h * w),
                                  cable_num_components=cable_n, tape_num_components=tape_n,
                                  tape_largest_area_ratio=(tape_largest / tape_px) if tape_n > 0 else 0.0,
                                  defect_areas=areas, total_defect_area=sum(areas.values())))
    return res[0] if single else res


def enhance_image(frame):
    """enhance_image (infer_enhanced_burr.py:38-66): BGR2LAB, CLAHE(3.0, 8x8) on L, LAB2BGR,
    fastNlMeansDenoisingColored(10, 10, 7, 21), 3x3 sharpen -- bit-exact with cv2 on uint8.
    frame: u8 BGR [H,W,3] (numpy -> numpy) or a CUDA tensor [H,W,3] / [B,H,W,3] (stays on the GPU).
    W and H must be multiples of 8 (the reference only ever passes 800x448 frames)."""
    f, was_np, single = _batched(frame, 3)
    c = _ctx(f.device, f.shape[0], tuple(f.shape[1:3]))
    with torch.cuda.device(f.device):
        out = c.enhance(f)
    return _unbatch(out, was_np, single)


def overlay_blend(frame, mask_cable, mask_tape, mask_burr, roi):
    """The four cv2.addWeighted steps of visualize_two_stage (infer_two_stage_burr.py:134-154): dim outside
    the ROI, tint cable / tape / burr -- one device pass, bit-exact with cv2.  numpy in -> numpy out; CUDA
    tensors ([H,W,3] / [B,H,W,3] and [H,W] / [B,H,W]) stay on the GPU."""
    f, was_np, single = _batched(frame, 3)
    ms = [_batched(m, 2)[0].to(f.device) for m in (mask_cable, mask_tape, mask_burr)]
    c = _ctx(f.device, f.shape[0], tuple(f.shape[1:3]))
    with torch.cuda.device(f.device):
        out = c.overlay(f, ms[0], ms[1], ms[2], roi)
    return _unbatch(out, was_np, single)


_roi_layers = {}        # (H, W, roi) -> u8 [H,W] plane holding the ROI box and its label as cv2 rasterises them


def _roi_layer(h, w, roi):
    """cv2.rectangle(.., 2) + cv2.putText('ROI', .., FONT_HERSHEY_SIMPLEX, 0.6, .., 2) of infer_two_stage_burr.py:156-158 on a
    blank plane (1 where they paint): both write one solid colour, so painting that colour where the plane is set equals
    drawing them on the frame.  Rasterised once per (frame size, ROI) on the host -- a polyline and three Hershey glyphs."""
    import cv2
    key = (h, w, tuple(int(v) for v in roi))
    if key not in _roi_layers:
        x1, y1, x2, y2 = key[2]
        layer = np.zeros((h, w), np.uint8)
        cv2.rectangle(layer, (x1, y1), (x2, y2), 1, 2)
        cv2.putText(layer, "ROI", (x1 + 5, y1 + 25), cv2.FONT_HERSHEY_SIMPLEX, 0.6, 1, 2)
        if len(_roi_layers) > 16:
            _roi_layers.clear()
        _roi_layers[key] = layer
    return _roi_layers[key]


def annotate_on_device(c, f, mask_cable, mask_tape, mask_burr, roi, draw_roi_box=True):
    """The vector half of visualize_two_stage (infer_two_stage_burr.py:156-166) painted in place on the CUDA frames f u8
    [B,H,W,3] through context c: ROI box + label (host-rasterised once per ROI, ub_paint_layer), then the cable, tape
    (thickness 2) and burr (thickness 3) contours (ub_draw_contours); masks u8 [B,H,W] on the same device.
    Returns the int32 [B] redo flags of the burr contours (frames where cv2 must redraw them on top, see
    ub_draw_contours), or None -- nothing drawn -- when a frame is larger than the per-frame flood holds."""
    b, h, w, _ = f.shape
    if 2 * (h + 2) * ((w + 31) // 32 + 2) * 4 > 200 * 1024:
        return None
    if draw_roi_box:
        c.paint_layer(f, torch.from_numpy(_roi_layer(h, w, roi)).to(f.device), [(0, 255, 255)])
    c.draw_contours(f, mask_cable.contiguous(), (0, 255, 0))
    c.draw_contours(f, mask_tape.contiguous(), (255, 0, 0))
    _, redo = c.draw_contours(f, mask_burr.contiguous(), (255, 0, 255), thickness=3)
    return redo


def draw_host_contours(img, mask_burr, mask_cable=None, mask_tape=None, roi=None):
    """cv2's share: the burr contours (thickness 3, infer_two_stage_burr.py:165-166) of a frame whose redo flag is set --
    and, for frames the device path does not take, the box and the other two contour sets before them.
    img numpy [H,W,3], drawn in place."""
    import cv2
    todo = []
    if mask_cable is not None:
        if roi is not None:
            x1, y1, x2, y2 = roi
            cv2.rectangle(img, (x1, y1), (x2, y2), (0, 255, 255), 2)
            cv2.putText(img, "ROI", (x1 + 5, y1 + 25), cv2.FONT_HERSHEY_SIMPLEX, 0.6, (0, 255, 255), 2)
        todo = [(mask_cable, (0, 255, 0), 2), (mask_tape, (255, 0, 0), 2)]
    for mask, colour, thick in todo + [(mask_burr, (255, 0, 255), 3)]:
        contours, _ = cv2.findContours(np.ascontiguousarray(mask), cv2.RETR_EXTERNAL, cv2.CHAIN_APPROX_SIMPLE)
        cv2.drawContours(img, contours, -1, colour, thick)
    return img


def draw_annotations(img, mask_cable, mask_tape, mask_burr, roi, draw_roi_box=True):
    """The vector part of visualize_two_stage (infer_two_stage_burr.py:156-170) over the blended frame(s), on the device
    (annotate_on_device).  cv2 only redraws the burr contours of frames whose mask comes within a pixel of the image border
    (OpenCV's odd-thickness rasteriser paints extra pixels there), and draws everything for frames too large for the
    device path.  img: numpy or CUDA tensor, [H,W,3] / [B,H,W,3]; masks likewise [H,W] / [B,H,W].  Returns numpy."""
    f, _, single = _batched(img, 3)
    mc, mt, mb = (_batched(m, 2)[0].to(f.device) for m in (mask_cable, mask_tape, mask_burr))
    b, h, w, _ = f.shape
    with torch.cuda.device(f.device):
        redo = annotate_on_device(_ctx(f.device, b, (h, w)), f, mc, mt, mb, roi, draw_roi_box)
        out = f.cpu().numpy()
        redo = redo.cpu().numpy() if redo is not None else None
    if redo is None or redo.any():
        host = [(m.cpu().numpy() if isinstance(m, torch.Tensor) else np.asarray(m)) for m in (mask_burr, mask_cable, mask_tape)]
        host = [m[None] if m.ndim == 2 else m for m in host]
        for i in range(b):
            if redo is None:
                draw_host_contours(out[i], host[0][i], host[1][i], host[2][i], roi if draw_roi_box else None)
            elif redo[i]:
                draw_host_contours(out[i], host[0][i])
    return out[0] if single else out


def visualize_two_stage(frame, mask_cable, mask_tape, mask_burr, roi_orig, draw_roi_box=True):
    """Drop-in for visualize_two_stage(frame, mask_cable, mask_tape, mask_burr, roi_orig, draw_roi_box)
    (infer_two_stage_burr.py:130-170; visualize_detection of infer_enhanced_burr.py:150-192 with the ROI as a
    tuple): numpy in, numpy out."""
    f, _, _ = _batched(frame, 3)
    out = overlay_blend(f, mask_cable, mask_tape, mask_burr, roi_orig)          # stays on the device
    return draw_annotations(out[0] if np.ndim(frame) == 3 else out, mask_cable, mask_tape, mask_burr, roi_orig, draw_roi_box)
