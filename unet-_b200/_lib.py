"""ctypes binding of libunetb200.so (the C ABI declared in include/unetb200.h).

The shared library is built in-tree by ``make -C unet-_b200/csrc`` (see
__graft_entry__.build).  There is no fallback: if the library is missing, or the device
is not sm_100, every entry point raises.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np
import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# UNETB200_LIB: an alternative build of the same library (tools/build_trace_lib.sh)
LIB_PATH = os.environ.get("UNETB200_LIB") or os.path.join(_HERE, "libunetb200.so")
_lib = None

VARIANT_STANDARD, VARIANT_ENHANCED, VARIANT_LAPLACIAN = 0, 1, 2
FLAG_ENHANCE_IMAGE = 0x100
CVT_BGR2LAB, CVT_LBGR2LAB, CVT_LAB2BGR, CVT_LAB2LBGR = 0, 1, 2, 3
T_CLASSES = ("conv3x3", "net_glue", "prepost", "stencil_morph", "ccl", "enhance")


class UbError(RuntimeError):
    pass


class NamedTensor(C.Structure):
    _fields_ = [("name", C.c_char_p), ("data_host", C.c_void_p), ("numel", C.c_int64)]


class Roi(C.Structure):
    _fields_ = [("x1", C.c_int32), ("y1", C.c_int32), ("x2", C.c_int32), ("y2", C.c_int32)]


class BurrParams(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("min_area", "max_area", "max_aspect", "min_side", "canny_low",
                                         "canny_high", "band_k", "close_k", "open_k")]


# name -> (restype, argtypes); must list every symbol include/unetb200.h declares
_P = C.c_void_p
_I = C.c_int
SIGNATURES = {
    "ub_create": (_I, [C.POINTER(_P), _I, _I, _I, _I, _I, _I, _I]),
    "ub_destroy": (None, [_P]),
    "ub_last_error": (C.c_char_p, [_P]),
    "ub_version": (_I, []),
    "ub_set_weights": (_I, [_P, C.POINTER(NamedTensor), _I]),
    "ub_forward": (_I, [_P, _P, _I, _I, _I, _P, _P]),
    "ub_debug_set_conv_impl": (_I, [_P, _I]),
    "ub_preprocess": (_I, [_P, _P, _I, _I, _I, _P, _P, _P]),
    "ub_normalize_frames": (_I, [_P, _P, _I, _I, _I, _I, _P, _I, _I, _P]),
    "ub_masks_from_logits": (_I, [_P, _P, _I, _I, _I, _I, _I, C.POINTER(Roi), _P, _P, _P, _P]),
    "ub_overlay": (_I, [_P, _P, _P, _P, _P, _I, _I, _I, C.POINTER(Roi), _P, _P]),
    "ub_draw_contours": (_I, [_P, _P, _P, _I, _I, _I, _I, _I, _I, _I, _P, _P]),
    "ub_paint_layer": (_I, [_P, _P, _P, _I, _I, _I, _I, _P, _P]),
    "ub_contour_stamp": (C.c_ulonglong, [_I, _I]),
    "ub_diameter_metrics": (_I, [_P, _P, _I, _I, _I, _I, _I, _I, _I, _P, _P]),
    "ub_analyze_defects": (_I, [_P, _P, _I, _I, _I, _I, _I, _I, _P, _P]),
    "ub_width_profiles": (_I, [_P, _P, _I, _I, _I, _I, _I, _P, _P]),
    "ub_filter_cable_by_shape": (_I, [_P, _P, _I, _I, _I, _I, C.c_double, C.c_double, _I, _P, _P]),
    "ub_constrain_tape_to_ring": (_I, [_P, _P, _P, _I, _I, _I, _I, _I, _P, _P]),
    "ub_gaussian5": (_I, [_P, _P, _I, _I, _I, _P, _P]),
    "ub_canny": (_I, [_P, _P, _I, _I, _I, _I, _I, _P, _P]),
    "ub_blur_canny": (_I, [_P, _P, _I, _I, _I, _I, _I, _P, _P]),
    "ub_morph": (_I, [_P, _I, _I, _P, _I, _I, _I, _I, _P, _P]),
    "ub_ccl": (_I, [_P, _P, _I, _I, _I, _P, _P, _I, _P, _P]),
    "ub_sobel_edges": (_I, [_P, _P, _I, _I, _I, _I, _P, _P]),
    "ub_sobel_edges_static": (_I, [_P, _I, _I, _I, _I, _P, _P]),
    "ub_laplacian_edges": (_I, [_P, _P, _I, _I, _I, _I, _P, _P]),
    "ub_burr": (_I, [_P, _I, _P, _P, _I, _I, _I, C.POINTER(BurrParams), _P, _P, _P]),
    "ub_cvt_lab": (_I, [_P, _I, _P, _I, _I, _I, _P, _P]),
    "ub_clahe": (_I, [_P, _P, _I, _I, _I, C.c_double, _I, _I, _P, _P]),
    "ub_nlm": (_I, [_P, _P, _I, _I, _I, _I, _P, _P]),
    "ub_sharpen3": (_I, [_P, _P, _I, _I, _I, _P, _P]),
    "ub_enhance": (_I, [_P, _P, _I, _I, _I, _P, _P]),
    "ub_pipeline": (_I, [_P, _P, _I, _I, C.POINTER(Roi), C.POINTER(BurrParams), _P, _P, _P]),
    "ub_pipeline_graph": (_I, [_P, _P, _I, _I, C.POINTER(Roi), C.POINTER(BurrParams), _P, _P, _P]),
    "ub_pipeline_host": (_I, [_P, _P, _I, _I, C.POINTER(Roi), C.POINTER(BurrParams), _P, _P]),
    "ub_pipeline_submit": (_I, [_P, _P, _I, _I, C.POINTER(Roi), C.POINTER(BurrParams), _P, _I]),
    "ub_pipeline_wait": (_I, [_P, _I]),
    "ub_pipeline_submit_src": (_I, [_P, _P, _I, _I, _I, _I, _I, C.POINTER(Roi), C.POINTER(BurrParams), _P, _I]),
    "ub_set_option": (_I, [_P, C.c_char_p, _I]),
    "ub_get_stat": (_I, [_P, C.c_char_p, C.POINTER(C.c_int64), _I]),
    "ub_debug_conv": (_I, [_P, _I, _P, _P, _I, _I, _I, _P, _P]),
    "ub_debug_conv_up": (_I, [_P, _I, _P, _P, _I, _I, _I, _I, _P, _P]),
    "ub_debug_activation": (_I, [_P, _I, _I, _I, _I, _I, _P, _P]),
    "ub_timing_enable": (_I, [_P, _I]),
    "ub_timing_read": (_I, [_P, C.POINTER(C.c_float), C.POINTER(C.c_int64), _I]),
    "ub_launch_count": (C.c_int64, [_P]),
}


def build(verbose: bool = False) -> str:
    """Compile libunetb200.so for sm_100a with nvcc (cross-compiles without a GPU)."""
    res = subprocess.run(["make", "-C", os.path.join(_HERE, "csrc"), "-j8"], capture_output=True, text=True)
    if verbose or res.returncode != 0:
        print(res.stdout[-4000:])
        print(res.stderr[-4000:])
    if res.returncode != 0:
        raise UbError("building libunetb200.so failed")
    return LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise UbError(f"{LIB_PATH} is missing: build it with `make -C unet-_b200/csrc` "
                          "(python -c 'import __graft_entry__ as g; g.build()'); there is no CPU fallback")
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def _ptr(t):
    if t is None:
        return None
    if isinstance(t, torch.Tensor):
        assert t.is_contiguous(), "tensor must be contiguous"
        return t.data_ptr()
    if isinstance(t, np.ndarray):
        assert t.flags["C_CONTIGUOUS"]
        return t.ctypes.data
    return int(t)


def _stream(device):
    return torch.cuda.current_stream(device).cuda_stream


def standard_params(min_area=30, max_area=800) -> BurrParams:
    """detect_burrs_on_cable constants (infer_two_stage_burr.py:78-117)."""
    return BurrParams(min_area, max_area, 5, 4, 50, 150, 8, 3, 2)


def enhanced_params(min_area=50, max_area=500) -> BurrParams:
    """detect_burrs_enhanced constants (infer_enhanced_burr.py:82-136)."""
    return BurrParams(min_area, max_area, 6, 5, 30, 100, 25, 5, 3)


def rulebased_params(band_out=10, laplacian_threshold=30, min_area=20, max_area=500) -> BurrParams:
    """get_burr_mask_rulebased / BurrConfig (src/refactor/burr_detector.py:11-66, src/refactor/config.py:98-103)."""
    return BurrParams(min_area, max_area, 1 << 20, 0, laplacian_threshold, 0, 2 * band_out + 1, 1, 1)


class Context:
    """Owns one ub_ctx (device scratch for up to max_batch frames)."""

    def __init__(self, device=0, max_batch=1, frame_hw=(448, 800), net_hw=(512, 512), num_classes=3):
        if not torch.cuda.is_available():
            raise UbError("no CUDA device: libunetb200 has no CPU fallback")
        self.device = torch.device("cuda", device if isinstance(device, int) else (device.index or 0))
        self.max_batch, self.frame_hw, self.net_hw, self.num_classes = max_batch, tuple(frame_hw), tuple(net_hw), num_classes
        self._h = C.c_void_p()
        self._check(lib().ub_create(C.byref(self._h), self.device.index, max_batch, frame_hw[0], frame_hw[1],
                                    net_hw[0], net_hw[1], num_classes), None)
        self._keep = None

    # -- plumbing
    def _check(self, code, h="self"):
        if code != 0:
            msg = lib().ub_last_error(self._h if h == "self" else None)
            raise UbError(f"libunetb200 error {code}: {msg.decode() if msg else '?'}")

    def close(self):
        if self._h:
            lib().ub_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _dev(self, t, dtype):
        assert t.is_cuda and t.dtype == dtype and t.is_contiguous(), (t.device, t.dtype)
        return t.data_ptr()

    @property
    def stream(self):
        return _stream(self.device)

    # -- stage 1
    def set_weights(self, state_dict):
        items, keep = [], []
        for k, v in state_dict.items():
            if k.endswith("num_batches_tracked"):
                continue
            a = v.detach().to("cpu", torch.float32).contiguous()
            keep.append(a)
            items.append((k.encode(), a))
        arr = (NamedTensor * len(items))()
        for i, (k, a) in enumerate(items):
            arr[i] = NamedTensor(k, a.data_ptr(), a.numel())
        self._check(lib().ub_set_weights(self._h, arr, len(items)))

    def set_conv_impl(self, impl: int):
        self._check(lib().ub_debug_set_conv_impl(self._h, impl))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        b, c, h, w = x.shape
        assert c == 3
        out = torch.empty((b, self.num_classes, h, w), dtype=torch.float32, device=x.device)
        self._check(lib().ub_forward(self._h, self._dev(x, torch.float32), b, h, w, out.data_ptr(), self.stream))
        return out

    # -- pre / post
    def preprocess(self, frames: torch.Tensor, want_x=True, want_gray=True):
        b, h, w, _ = frames.shape
        x = torch.empty((b, 3) + self.net_hw, dtype=torch.float32, device=frames.device) if want_x else None
        g = torch.empty((b, h, w), dtype=torch.uint8, device=frames.device) if want_gray else None
        self._check(lib().ub_preprocess(self._h, self._dev(frames, torch.uint8), b, h, w, _ptr(x), _ptr(g), self.stream))
        return x, g

    def normalize_frames(self, src: torch.Tensor, size_wh=(800, 448), rotate=False) -> torch.Tensor:
        b, sh, sw, _ = src.shape
        dst = torch.empty((b, size_wh[1], size_wh[0], 3), dtype=torch.uint8, device=src.device)
        self._check(lib().ub_normalize_frames(self._h, self._dev(src, torch.uint8), b, sh, sw, int(rotate),
                                              dst.data_ptr(), size_wh[1], size_wh[0], self.stream))
        return dst

    def masks_from_logits(self, logits: torch.Tensor, frame_hw, roi):
        b, c, nh, nw = logits.shape
        H, W = frame_hw
        pred = torch.empty((b, nh, nw), dtype=torch.uint8, device=logits.device)
        cable = torch.empty((b, H, W), dtype=torch.uint8, device=logits.device)
        tape = torch.empty_like(cable)
        r = Roi(*roi)
        self._check(lib().ub_masks_from_logits(self._h, self._dev(logits, torch.float32), b, nh, nw, H, W, C.byref(r),
                                               pred.data_ptr(), cable.data_ptr(), tape.data_ptr(), self.stream))
        return pred, cable, tape

    # -- stage 2 operators
    def overlay(self, frames, mask_cable, mask_tape, mask_burr, roi):
        """Blend chain of visualize_two_stage: frames u8 [B,H,W,3], masks u8 [B,H,W] -> u8 [B,H,W,3]."""
        b, h, w, _ = frames.shape
        out = torch.empty_like(frames)
        r = Roi(*[int(v) for v in roi])
        self._check(lib().ub_overlay(self._h, self._dev(frames, torch.uint8), self._dev(mask_cable, torch.uint8),
                                     self._dev(mask_tape, torch.uint8), self._dev(mask_burr, torch.uint8), b, h, w,
                                     C.byref(r), out.data_ptr(), self.stream))
        return out

    def draw_contours(self, frames, mask, bgr, thickness=2):
        """cv2.findContours(RETR_EXTERNAL, CHAIN_APPROX_SIMPLE) + cv2.drawContours(.., bgr, thickness) in place: frames u8
        [B,H,W,3], mask u8 [B,H,W].  thickness 2 -> frames; thickness 3 -> (frames, redo int32 [B]): frames whose flag is set
        must have their contours redrawn by cv2 on top (OpenCV paints extra pixels near the image border)."""
        b, h, w, _ = frames.shape
        redo = torch.empty((b,), dtype=torch.int32, device=frames.device) if thickness == 3 else None
        self._check(lib().ub_draw_contours(self._h, self._dev(frames, torch.uint8), self._dev(mask, torch.uint8), b, h, w,
                                           int(bgr[0]), int(bgr[1]), int(bgr[2]), int(thickness),
                                           redo.data_ptr() if redo is not None else None, self.stream))
        return frames if redo is None else (frames, redo)

    def paint_layer(self, frames, layer, colours):
        """frames[q] = colours[layer[q] - 1] where layer[q] != 0, in place; layer u8 [H,W] (shared) or [B,H,W]; up to three BGR colours."""
        b, h, w, _ = frames.shape
        c9 = (C.c_uint8 * 9)(*([int(v) for c in colours for v in c] + [0] * 9)[:9])
        self._check(lib().ub_paint_layer(self._h, self._dev(frames, torch.uint8), self._dev(layer, torch.uint8), int(layer.dim() == 3),
                                         b, h, w, c9, self.stream))
        return frames

    def diameter_metrics(self, pred, cable_cls=1, tape_cls=2, min_area=50, min_valid_rows=20):
        """pred u8 [B,H,W] class ids -> int32 [B,8] (dc_px / dt_px as float32 bit patterns, valid_rows, cable_area, tape_area)."""
        b, h, w = pred.shape
        out = torch.empty((b, 8), dtype=torch.int32, device=pred.device)
        self._check(lib().ub_diameter_metrics(self._h, self._dev(pred, torch.uint8), b, h, w, cable_cls, tape_cls, min_area,
                                              min_valid_rows, out.data_ptr(), self.stream))
        return out

    def analyze_defects(self, pred, cable_cls=1, tape_cls=2, hole_min_size=10):
        """pred u8 [B,H,W] class ids -> int32 [B,16] (see ub_analyze_defects)."""
        b, h, w = pred.shape
        out = torch.empty((b, 16), dtype=torch.int32, device=pred.device)
        self._check(lib().ub_analyze_defects(self._h, self._dev(pred, torch.uint8), b, h, w, cable_cls, tape_cls, hole_min_size,
                                             out.data_ptr(), self.stream))
        return out

    def width_profiles(self, pred, cable_cls=1, tape_cls=2):
        """pred u8 [B,H,W] -> float32 [B,2,H]: smoothed per-row widths of the cable and the tape class."""
        b, h, w = pred.shape
        out = torch.empty((b, 2, h), dtype=torch.float32, device=pred.device)
        self._check(lib().ub_width_profiles(self._h, self._dev(pred, torch.uint8), b, h, w, cable_cls, tape_cls, out.data_ptr(), self.stream))
        return out

    def filter_cable_by_shape(self, mask, min_area, min_aspect, max_center_offset, roi_width):
        b, h, w = mask.shape
        out = torch.empty_like(mask)
        self._check(lib().ub_filter_cable_by_shape(self._h, self._dev(mask, torch.uint8), b, h, w, int(min_area), float(min_aspect),
                                                   float(max_center_offset), int(roi_width), out.data_ptr(), self.stream))
        return out

    def constrain_tape_to_ring(self, tape, cable, dilate_k, erode_k):
        b, h, w = tape.shape
        out = torch.empty_like(tape)
        self._check(lib().ub_constrain_tape_to_ring(self._h, self._dev(tape, torch.uint8), self._dev(cable, torch.uint8), b, h, w,
                                                    int(dilate_k), int(erode_k), out.data_ptr(), self.stream))
        return out

    def _u8op(self, fn, src, *args):
        b, h, w = src.shape
        dst = torch.empty_like(src)
        self._check(fn(self._h, self._dev(src, torch.uint8), b, h, w, *args, dst.data_ptr(), self.stream))
        return dst

    def gaussian5(self, src):
        return self._u8op(lib().ub_gaussian5, src)

    def canny(self, src, low, high):
        return self._u8op(lib().ub_canny, src, low, high)

    def blur_canny(self, gray, low, high):
        return self._u8op(lib().ub_blur_canny, gray, low, high)

    def sobel_edges(self, gray, thresh=50):
        return self._u8op(lib().ub_sobel_edges, gray, thresh)

    def laplacian_edges(self, gray, thresh=15):
        return self._u8op(lib().ub_laplacian_edges, gray, thresh)

    def morph(self, op, k, src, value=1):
        b, h, w = src.shape
        dst = torch.empty_like(src)
        code = {"dilate": 0, "erode": 1, "close": 2, "open": 3}[op]
        self._check(lib().ub_morph(self._h, code, k, self._dev(src, torch.uint8), b, h, w, value, dst.data_ptr(), self.stream))
        return dst

    def ccl(self, src, max_labels=4096):
        b, h, w = src.shape
        labels = torch.empty((b, h, w), dtype=torch.int32, device=src.device)
        stats = torch.empty((b, max_labels, 5), dtype=torch.int32, device=src.device)
        n = torch.empty((b,), dtype=torch.int32, device=src.device)
        self._check(lib().ub_ccl(self._h, self._dev(src, torch.uint8), b, h, w, labels.data_ptr(), stats.data_ptr(),
                                 max_labels, n.data_ptr(), self.stream))
        return n, labels, stats

    def burr(self, gray, mask, variant=VARIANT_STANDARD, params=None):
        b, h, w = gray.shape
        out = torch.empty_like(gray)
        px = torch.empty((b,), dtype=torch.int32, device=gray.device)
        p = params if params is not None else (standard_params(), enhanced_params(), rulebased_params())[variant]
        self._check(lib().ub_burr(self._h, variant, self._dev(gray, torch.uint8), self._dev(mask, torch.uint8), b, h, w,
                                  C.byref(p), out.data_ptr(), px.data_ptr(), self.stream))
        return out, px

    # -- enhance_image pieces (u8 [B,H,W,3] unless noted)
    def cvt_lab(self, src, code):
        b, h, w, _ = src.shape
        dst = torch.empty_like(src)
        self._check(lib().ub_cvt_lab(self._h, code, self._dev(src, torch.uint8), b, h, w, dst.data_ptr(), self.stream))
        return dst

    def clahe(self, src, clip_limit=3.0, tiles=(8, 8)):
        """src u8 [B,H,W]"""
        b, h, w = src.shape
        dst = torch.empty_like(src)
        self._check(lib().ub_clahe(self._h, self._dev(src, torch.uint8), b, h, w, float(clip_limit), tiles[0], tiles[1],
                                   dst.data_ptr(), self.stream))
        return dst

    def nlm(self, src):
        """src u8 [B,H,W] or [B,H,W,2]; fastNlMeansDenoising(h=10, 7, 21)"""
        b, h, w = src.shape[:3]
        cn = 1 if src.dim() == 3 else src.shape[3]
        dst = torch.empty_like(src)
        self._check(lib().ub_nlm(self._h, self._dev(src, torch.uint8), b, h, w, cn, dst.data_ptr(), self.stream))
        return dst

    def sharpen3(self, src):
        b, h, w, _ = src.shape
        dst = torch.empty_like(src)
        self._check(lib().ub_sharpen3(self._h, self._dev(src, torch.uint8), b, h, w, dst.data_ptr(), self.stream))
        return dst

    def enhance(self, frames):
        b, h, w, _ = frames.shape
        dst = torch.empty_like(frames)
        self._check(lib().ub_enhance(self._h, self._dev(frames, torch.uint8), b, h, w, dst.data_ptr(), self.stream))
        return dst

    # -- whole step
    def pipeline(self, frames: torch.Tensor, variant=VARIANT_STANDARD, roi=None, params=None, want_masks=False,
                 stats=None, masks=None, graph=False):
        """graph=True replays the step as a CUDA graph (ub_pipeline_graph): pass the same frames / stats / masks tensors
        every time, the graph is keyed on their addresses."""
        b = frames.shape[0]
        assert tuple(frames.shape[1:3]) == self.frame_hw
        if stats is None:
            stats = torch.empty((b, 4), dtype=torch.int32, device=frames.device)
        if masks is None and want_masks:
            masks = torch.empty((b, 3) + self.frame_hw, dtype=torch.uint8, device=frames.device)
        r = C.byref(Roi(*roi)) if roi is not None else None
        p = C.byref(params) if params is not None else None
        fn = lib().ub_pipeline_graph if graph else lib().ub_pipeline
        self._check(fn(self._h, self._dev(frames, torch.uint8), b, variant, r, p, stats.data_ptr(), _ptr(masks), self.stream))
        return stats, masks

    def pipeline_host(self, frames, stats_out, variant=VARIANT_STANDARD, roi=None, params=None, masks_out=None):
        """frames / stats_out: pinned host tensors or numpy arrays.  Synchronous."""
        b = frames.shape[0]
        r = C.byref(Roi(*roi)) if roi is not None else None
        p = C.byref(params) if params is not None else None
        self._check(lib().ub_pipeline_host(self._h, _ptr(frames), b, variant, r, p, _ptr(stats_out), _ptr(masks_out)))
        return stats_out

    def pipeline_submit(self, frames, stats_out, slot, variant=VARIANT_STANDARD, roi=None, params=None):
        """Asynchronous pipeline_host into `slot` (0/1); frames / stats_out must be pinned and stay alive until pipeline_wait."""
        r = C.byref(Roi(*roi)) if roi is not None else None
        p = C.byref(params) if params is not None else None
        self._check(lib().ub_pipeline_submit(self._h, _ptr(frames), frames.shape[0], variant, r, p, _ptr(stats_out), slot))

    def pipeline_submit_src(self, src, stats_out, slot, rotate, variant=VARIANT_STANDARD, roi=None, params=None):
        """pipeline_submit for frames at the source resolution: src pinned uint8 [b,sh,sw,3]; rotate + resize run on the device."""
        r = C.byref(Roi(*roi)) if roi is not None else None
        p = C.byref(params) if params is not None else None
        b, sh, sw, _ = src.shape
        self._check(lib().ub_pipeline_submit_src(self._h, _ptr(src), b, sh, sw, int(rotate), variant, r, p, _ptr(stats_out), slot))

    def set_option(self, name: str, value: int):
        self._check(lib().ub_set_option(self._h, name.encode(), int(value)))

    def get_stat(self, name: str, reset=False) -> int:
        v = C.c_int64(0)
        self._check(lib().ub_get_stat(self._h, name.encode(), C.byref(v), int(reset)))
        return int(v.value)

    def pipeline_wait(self, slot):
        self._check(lib().ub_pipeline_wait(self._h, slot))

    # -- test hooks
    def debug_conv(self, layer, x0, x1, cout):
        b, _, h, w = x0.shape
        out = torch.empty((b, cout, h, w), dtype=torch.float32, device=x0.device)
        self._check(lib().ub_debug_conv(self._h, layer, self._dev(x0, torch.float32),
                                        self._dev(x1, torch.float32) if x1 is not None else None, b, h, w,
                                        out.data_ptr(), self.stream))
        return out

    def debug_conv_up(self, layer, x0, x1_low, cout, fused):
        b, _, h, w = x0.shape
        out = torch.empty((b, cout, h, w), dtype=torch.float32, device=x0.device)
        self._check(lib().ub_debug_conv_up(self._h, layer, self._dev(x0, torch.float32), self._dev(x1_low, torch.float32),
                                           b, h, w, int(fused), out.data_ptr(), self.stream))
        return out

    def debug_activation(self, index, b, c, h, w):
        out = torch.empty((b, c, h, w), dtype=torch.float32, device=self.device)
        self._check(lib().ub_debug_activation(self._h, index, b, c, h, w, out.data_ptr(), self.stream))
        return out

    # -- instrumentation
    def timing_enable(self, on=True):
        self._check(lib().ub_timing_enable(self._h, int(on)))

    def timing_read(self, reset=True):
        ms = (C.c_float * len(T_CLASSES))()
        n = (C.c_int64 * len(T_CLASSES))()
        self._check(lib().ub_timing_read(self._h, ms, n, int(reset)))
        return {k: (ms[i], n[i]) for i, k in enumerate(T_CLASSES)}

    def launch_count(self) -> int:
        return int(lib().ub_launch_count(self._h))
