"""oracle/cvlab.py -- TEST INFRASTRUCTURE ONLY (never imported by the product path).

Numpy restatement of the OpenCV routines that `enhance_image`
(infer_enhanced_burr.py:38-66 of the reference) calls besides CLAHE / filter2D
(those two are in oracle/cvnp.py):

  bgr2lab_u8 / lab2bgr_u8     cv2.cvtColor(COLOR_BGR2LAB / COLOR_LAB2BGR), :46, :55, and the
                              linear-gamma variants COLOR_LBGR2Lab / COLOR_Lab2LBGR that
                              fastNlMeansDenoisingColored uses internally
  nlm_u8                      cv2.fastNlMeansDenoising on 1- and 2-channel uint8 images
  nlm_colored                 cv2.fastNlMeansDenoisingColored(img, None, 10, 10, 7, 21), :58

The algorithm lives in a third-party dependency of the reference (opencv-python, here
4.13.0: imgproc/src/color_lab.cpp, photo/src/denoising.cpp,
fast_nlmeans_denoising_invoker.hpp); what is restated is its published integer algorithm.
Pinned by tests/test_oracle_cvlab.py: the colour conversions are compared with cv2 on ALL
2^24 inputs, the denoiser on seeded images including ragged sizes.
"""
from __future__ import annotations

import numpy as np

# ---------------------------------------------------------------------------
# 8-bit BGR <-> CIE L*a*b* (fixed point, table driven)
# ---------------------------------------------------------------------------
LAB_SHIFT = 12
GAMMA_SHIFT = 3
LAB_SHIFT2 = LAB_SHIFT + GAMMA_SHIFT
CBRT_TAB_SIZE_B = 256 * 3 // 2 * (1 << GAMMA_SHIFT)
BASE_SHIFT = 14
BASE = 1 << BASE_SHIFT
INV_GAMMA_TAB_SIZE = 1 << 12
MIN_AB = -8145
AB_TAB_SIZE = BASE * 9 // 4

_RGB2XYZ = np.array([0.412453, 0.357580, 0.180423,
                     0.212671, 0.715160, 0.072169,
                     0.019334, 0.119193, 0.950227], np.float64)
_XYZ2RGB = np.array([3.240479, -1.53715, -0.498535,
                     -0.969256, 1.875991, 0.041556,
                     0.055648, -0.204043, 1.057311], np.float64)
_D65 = np.array([0.950456, 1.0, 1.088754], np.float64)


def _cvround(x):
    """cvRound: round half to even (lrint)."""
    return np.rint(x).astype(np.int64)


def soft_cbrt(x):
    """cv::cbrt(softfloat) (core/src/softfloat.cpp): the quartic rational approximation of
    cv::cubeRoot evaluated in binary64 on the mantissa scaled into [0.125, 1), whose
    binary64 fraction is then TRUNCATED to 23 bits (not rounded)."""
    x = np.asarray(x, np.float32)
    ix = x.view(np.int32) & 0x7fffffff
    ex = (ix >> 23) - 127
    shx = np.fmod(ex, 3)                                  # C remainder
    shx = shx - np.where(shx >= 0, 3, 0)
    ex3 = (ex - shx) // 3
    fr = ((ix & ((1 << 23) - 1)) | ((shx + 127) << 23)).astype(np.int32).view(np.float32).astype(np.float64)
    num = ((((45.2548339756803022511987494 * fr + 192.2798368355061050458134625) * fr
             + 119.1654824285581628956914143) * fr + 13.43250139086239872172837314) * fr
           + 0.1636161226585754240958355063)
    den = ((((14.80884093219134573786480845 * fr + 151.9714051044435648658557668) * fr
             + 168.5254414101568283957668343) * fr + 33.9905941350215598754191872) * fr + 1.0)
    q = (num / den).view(np.int64) & ~np.int64((1 << 29) - 1)
    r = q.view(np.float64).astype(np.float32)
    out = (r.view(np.int32) + (ex3 << 23).astype(np.int32)).view(np.float32)
    return np.where(x == 0, np.float32(0), out)


_TABS = None


def lab_tables():
    """The lookup tables of color_lab.cpp (initLabTabs).  OpenCV evaluates them with its
    softfloat (binary32) type; +,*,/ there are correctly rounded, so numpy float32 is
    the same arithmetic.  pow / cbrt go through float64 here and are rounded to integers;
    the exhaustive test shows the integer tables agree."""
    global _TABS
    if _TABS is not None:
        return _TABS
    f32 = np.float32
    i256 = np.arange(256)
    x = (i256.astype(f32) / f32(255.0)).astype(np.float64)
    g = np.where(x <= 0.04045, x / 12.92, np.power((x + 0.055) / 1.055, 2.4))
    srgb_gamma = _cvround(f32(255 * (1 << GAMMA_SHIFT)) * g.astype(f32))
    lin_gamma = i256 * (1 << GAMMA_SHIFT)

    xi = (f32(1.0) / f32(INV_GAMMA_TAB_SIZE)) * np.arange(INV_GAMMA_TAB_SIZE).astype(f32)
    xd = xi.astype(np.float64)
    ig = np.where(xd <= 0.0031308, xd * 12.92, 1.055 * np.power(xd, 1.0 / 2.4) - 0.055)
    srgb_inv = _cvround(f32(255.0) * ig.astype(f32))
    lin_inv = np.trunc(f32(255.0) * xi).astype(np.int64)

    scale = f32(1.0) / (f32(255.0) * f32(1 << GAMMA_SHIFT))
    xc = scale * np.arange(CBRT_TAB_SIZE_B).astype(f32)
    lthresh = f32(216) / f32(24389)
    lscale = f32(841) / f32(108)
    lbias = f32(16) / f32(116)
    lin = (xc.astype(np.float64) * np.float64(lscale) + np.float64(lbias)).astype(f32)   # fused mulAdd
    cb = soft_cbrt(xc)
    cbrt_tab = _cvround(f32(1 << LAB_SHIFT2) * np.where(xc < lthresh, lin, cb))

    ytab = np.zeros(256, np.int64)
    fytab = np.zeros(256, np.int64)
    for i in range(256):
        if i <= 20:
            ytab[i] = _cvround(f32(i * BASE * 20 * 9) / f32(17 * 29 * 29 * 29))
            fytab[i] = _cvround(f32(BASE) * (f32(16) / f32(116) + f32(i * 5) / f32(3 * 17 * 29)))
        else:
            fy = f32(i * 100 * BASE) / f32(255 * 116) + f32(16 * BASE) / f32(116)
            fytab[i] = _cvround(fy)
            ytab[i] = _cvround(fy * fy * fy / f32(BASE * BASE))

    i = np.arange(MIN_AB, MIN_AB + AB_TAB_SIZE, dtype=np.int64)
    # C integer division truncates toward zero
    lo = np.trunc(i * 108 / 841).astype(np.int64) - (BASE * 16 // 116 * 108 // 841)
    hi = (i * i // BASE) * i // BASE
    ab_tab = np.where(i <= 3390, lo, hi)

    lshift = float(1 << LAB_SHIFT)
    fwd = np.zeros(9, np.int64)   # applied to (B, G, R) in memory order
    for r in range(3):
        c = _RGB2XYZ[r * 3:r * 3 + 3]
        fwd[r * 3 + 2] = _cvround(lshift * c[0] / _D65[r])
        fwd[r * 3 + 1] = _cvround(lshift * c[1] / _D65[r])
        fwd[r * 3 + 0] = _cvround(lshift * c[2] / _D65[r])
    inv = np.zeros(9, np.int64)   # rows produce (B, G, R) in memory order
    for r in range(3):            # r = x, y, z column
        c = _XYZ2RGB[[r, r + 3, r + 6]]
        inv[r + 0 * 3] = _cvround(lshift * c[2] * _D65[r])     # blue row
        inv[r + 3] = _cvround(lshift * c[1] * _D65[r])
        inv[r + 2 * 3] = _cvround(lshift * c[0] * _D65[r])     # red row
    _TABS = dict(srgb_gamma=srgb_gamma, lin_gamma=lin_gamma, srgb_inv=srgb_inv, lin_inv=lin_inv,
                 cbrt=cbrt_tab, y=ytab, fy=fytab, ab=ab_tab, fwd=fwd, inv=inv)
    return _TABS


def _descale(x, n):
    return (x + (1 << (n - 1))) >> n


def bgr2lab_u8(bgr: np.ndarray, srgb: bool = True) -> np.ndarray:
    """cvtColor(COLOR_BGR2LAB) (srgb) / COLOR_LBGR2Lab (linear) on uint8: RGB2Lab_b."""
    t = lab_tables()
    tab = t["srgb_gamma"] if srgb else t["lin_gamma"]
    c = t["fwd"]
    p0, p1, p2 = (tab[bgr[..., k]] for k in range(3))
    fx = t["cbrt"][_descale(p0 * c[0] + p1 * c[1] + p2 * c[2], LAB_SHIFT)]
    fy = t["cbrt"][_descale(p0 * c[3] + p1 * c[4] + p2 * c[5], LAB_SHIFT)]
    fz = t["cbrt"][_descale(p0 * c[6] + p1 * c[7] + p2 * c[8], LAB_SHIFT)]
    lscale = (116 * 255 + 50) // 100
    lshift = -((16 * 255 * (1 << LAB_SHIFT2) + 50) // 100)
    L = _descale(lscale * fy + lshift, LAB_SHIFT2)
    a = _descale(500 * (fx - fy) + 128 * (1 << LAB_SHIFT2), LAB_SHIFT2)
    b = _descale(200 * (fy - fz) + 128 * (1 << LAB_SHIFT2), LAB_SHIFT2)
    return np.clip(np.stack([L, a, b], -1), 0, 255).astype(np.uint8)


def lab2bgr_u8(lab: np.ndarray, srgb: bool = True) -> np.ndarray:
    """cvtColor(COLOR_LAB2BGR) (srgb) / COLOR_Lab2LBGR (linear) on uint8: Lab2RGBinteger."""
    t = lab_tables()
    tab = t["srgb_inv"] if srgb else t["lin_inv"]
    c = t["inv"]
    L = lab[..., 0].astype(np.int64)
    aa = lab[..., 1].astype(np.int64)
    bb = lab[..., 2].astype(np.int64)
    y = t["y"][L]
    ify = t["fy"][L]
    adiv = ((5 * aa * 53687 + (1 << 7)) >> 13) - 128 * BASE // 500
    bdiv = ((bb * 41943 + (1 << 4)) >> 9) - 128 * BASE // 200 + 1
    x = t["ab"][ify + adiv - MIN_AB]
    z = t["ab"][ify - bdiv - MIN_AB]
    shift = LAB_SHIFT + (BASE_SHIFT - 12)
    out = []
    for r in range(3):
        v = _descale(c[r * 3] * x + c[r * 3 + 1] * y + c[r * 3 + 2] * z, shift)
        out.append(tab[np.clip(v, 0, INV_GAMMA_TAB_SIZE - 1)])
    return np.clip(np.stack(out, -1), 0, 255).astype(np.uint8)


# ---------------------------------------------------------------------------
# Non-local means (photo module), uint8, squared-L2 distance
# ---------------------------------------------------------------------------

def nlm_weight_table(h: float, cn: int, tws: int = 7, sws: int = 21):
    """almost_dist2weight_ of FastNlMeansDenoisingInvoker (fast_nlmeans_denoising_invoker.hpp
    ctor): fixed-point weights indexed by (sum of squared differences over the template)
    >> bin_shift.  Returns (table int64[], bin_shift, fixed_point_mult)."""
    int_max = 2 ** 31 - 1
    max_estimate_sum = sws * sws * 255
    fixed_point_mult = int_max // max_estimate_sum          # min(., sampleMax<int>) never binds: 19096
    tsq = tws * tws
    shift = 0
    while (1 << shift) < tsq:
        shift += 1
    mult = float(1 << shift) / tsq
    max_dist = 255 * 255 * cn
    n = int(max_dist / mult + 1)
    d = np.arange(n, dtype=np.float64) * mult
    hh = np.float32(h)
    w = np.exp(-d / float(hh * hh * np.float32(cn)))
    tab = np.rint(fixed_point_mult * w).astype(np.int64)
    tab[tab < 0.001 * fixed_point_mult] = 0
    return tab, shift, fixed_point_mult


def nlm_u8(img: np.ndarray, h: float, tws: int = 7, sws: int = 21) -> np.ndarray:
    """cv2.fastNlMeansDenoising(img, None, h, tws, sws) for uint8 [H,W] or [H,W,2].
    photo/src/denoising.cpp fastNlMeansDenoising -> FastNlMeansDenoisingInvoker<uchar|Vec2b,
    int, unsigned, DistSquared, int>: REFLECT_101 border of sws/2 + tws/2, for every search
    offset the template SSD (all channels summed), weight = table[ssd >> shift],
    out = (sum w*p + sum w / 2) / sum w."""
    src = img if img.ndim == 3 else img[..., None]
    H, W, cn = src.shape
    tab, shift, _ = nlm_weight_table(h, cn, tws, sws)
    tr, sr = tws // 2, sws // 2
    bs = tr + sr
    ext = np.pad(src.astype(np.int64), ((bs, bs), (bs, bs), (0, 0)), mode="reflect")
    # centre window enlarged by the template radius
    c0 = ext[sr:sr + H + 2 * tr, sr:sr + W + 2 * tr]
    est = np.zeros((H, W, cn), np.int64)
    wsum = np.zeros((H, W), np.int64)
    for dy in range(-sr, sr + 1):
        for dx in range(-sr, sr + 1):
            sh = ext[sr + dy:sr + dy + H + 2 * tr, sr + dx:sr + dx + W + 2 * tr]
            d2 = ((c0 - sh) ** 2).sum(-1)
            ii = np.zeros((d2.shape[0] + 1, d2.shape[1] + 1), np.int64)
            ii[1:, 1:] = d2.cumsum(0).cumsum(1)
            ssd = ii[tws:, tws:] - ii[:-tws, tws:] - ii[tws:, :-tws] + ii[:-tws, :-tws]
            w = tab[np.minimum(ssd >> shift, len(tab) - 1)]
            est += w[..., None] * sh[tr:tr + H, tr:tr + W]
            wsum += w
    out = (est + (wsum // 2)[..., None]) // wsum[..., None]
    out = np.clip(out, 0, 255).astype(np.uint8)
    return out if img.ndim == 3 else out[..., 0]


def nlm_colored(bgr: np.ndarray, h: float = 10, h_color: float = 10, tws: int = 7, sws: int = 21) -> np.ndarray:
    """cv2.fastNlMeansDenoisingColored (photo/src/denoising.cpp): LINEAR-gamma BGR->Lab,
    NLM on L with h, NLM on the 2-channel (a,b) plane with h_color, Lab->linear BGR."""
    lab = bgr2lab_u8(bgr, srgb=False)
    L = nlm_u8(np.ascontiguousarray(lab[..., 0]), h, tws, sws)
    ab = nlm_u8(np.ascontiguousarray(lab[..., 1:]), h_color, tws, sws)
    return lab2bgr_u8(np.concatenate([L[..., None], ab], -1), srgb=False)


def enhance_stages(frame_bgr: np.ndarray) -> dict:
    """enhance_image (infer_enhanced_burr.py:38-66) from the restatements only, with every
    intermediate: lab, l_clahe, bgr_clahe, lab_lin, l_den, ab_den, denoised, out."""
    from . import cvnp
    lab = bgr2lab_u8(frame_bgr, srgb=True)                                  # :46
    l2 = cvnp.clahe_u8(np.ascontiguousarray(lab[..., 0]), 3.0, (8, 8))      # :50-51
    bgr = lab2bgr_u8(np.concatenate([l2[..., None], lab[..., 1:]], -1), srgb=True)   # :54-55
    lab_lin = bgr2lab_u8(bgr, srgb=False)                                   # :58 (inside cv2)
    l_den = nlm_u8(np.ascontiguousarray(lab_lin[..., 0]), 10.0)
    ab_den = nlm_u8(np.ascontiguousarray(lab_lin[..., 1:]), 10.0)
    den = lab2bgr_u8(np.concatenate([l_den[..., None], ab_den], -1), srgb=False)
    out = cvnp.sharpen3(den)                                                # :61-64
    return dict(lab=lab, l_clahe=l2, bgr_clahe=bgr, lab_lin=lab_lin, l_den=l_den, ab_den=ab_den,
                denoised=den, out=out)


def enhance(frame_bgr: np.ndarray) -> np.ndarray:
    return enhance_stages(frame_bgr)["out"]
