"""Times the stage-2 / pre-post entry points one by one through the C-ABI (CUDA events, planes resident in
HBM) and reports achieved GB/s on each call's own compulsory bytes (1 B/px per u8 plane read or
written).  B = 512 frames (183 MB per plane, larger than the 126 MB L2) unless given.
Usage: python tools/bench_stage2.py [frames] [reps]        (bench.py imports measure() for its `hbm_kernels` key)"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def measure(B=512, reps=10, log=None):
    import numpy as np
    import torch

    import unet_b200.synth as S
    from unet_b200 import _lib

    H, W = 448, 800
    P = H * W
    ctx = _lib.Context(torch.cuda.current_device(), max_batch=B, frame_hw=(H, W), net_hw=(16, 16))   # stage 2 only: tiny network planes
    base = np.concatenate([S.frames_cable(12, H, W, seed=1), S.frames_noise(4, H, W, seed=2)])
    frames3 = torch.from_numpy(np.concatenate([base] * (B // 16))).cuda()
    _, gray = ctx.preprocess(frames3, want_x=False, want_gray=True)
    mask = torch.from_numpy(np.stack([S.mask_cable_synth(H, W)] * B)).cuda()
    mask_t = torch.from_numpy(np.stack([S.mask_blobs(H, W, seed=3, fill=0.02)] * B)).cuda()
    mask_b = torch.from_numpy(np.stack([S.mask_blobs(H, W, seed=4, fill=0.004)] * B)).cuda()
    preds = torch.from_numpy(np.stack([S.pred_classes(H, W, seed=i) for i in range(8)] * (B // 8))).cuda()
    cand = ctx.blur_canny(gray, 50, 150)            # a realistic binary plane for the labelling row (Canny edges)
    roi = (218, 0, 421, 448)
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    peak = json.load(open(pk))["hbm_gbs"] if os.path.exists(pk) else 6549.8

    def timed(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(torch.cuda.current_stream())
        for _ in range(reps):
            fn()
        e1.record(torch.cuda.current_stream())
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    rows = [
        ("ub_overlay: 4 x addWeighted blend (LUT kernel)", lambda: ctx.overlay(frames3, mask, mask_t, mask_b, roi), 9 * P),
        ("ub_preprocess gray only: BGR2GRAY", lambda: ctx.preprocess(frames3, want_x=False, want_gray=True), 4 * P),
        ("ub_gaussian5: GaussianBlur 5x5 (k_blur5_cols)", lambda: ctx.gaussian5(gray), 2 * P),
        ("ub_blur_canny: k_edge_fused<1> + flood hysteresis (k_hyst_frame)", lambda: ctx.blur_canny(gray, 50, 150), 2 * P),
        ("ub_burr: edges + hysteresis + morphology chain (k_burr_chain_cols) + CCL + filter", lambda: ctx.burr(gray, mask), 3 * P),
        ("ub_ccl: connectedComponentsWithStats, labels in OpenCV order + stats", lambda: ctx.ccl(cand), 5 * P),
        ("ub_diameter_metrics: 2 x (CCL + largest component + row extents) + smoothing + medians", lambda: ctx.diameter_metrics(preds), P),
        ("ub_analyze_defects: 3 x CCL + CLOSE 5x5 + class histogram", lambda: ctx.analyze_defects(preds), P),
        ("ub_morph close 3x3 (k_morph_cols)", lambda: ctx.morph("close", 3, mask), 2 * P),
        ("ub_morph dilate 8x8 (k_morph_cols)", lambda: ctx.morph("dilate", 8, mask), 2 * P),
    ]
    out = {"frames": B, "hw": [H, W], "hbm_peak_gbs": peak, "peak_source": "MEASURED_PEAKS.json hbm_gbs" if os.path.exists(pk) else "fallback",
           "peak_note": "hbm_gbs is a measured COPY bandwidth (1 read : 1 write); a read-heavy kernel (BGR2GRAY: 3 : 1) can exceed it",
           "edge_fused": os.environ.get("UB_EDGE_FUSED", "1"),
           "env": {k: os.environ[k] for k in ("UB_BLUR_COLS", "UB_MORPH_WORDS", "UB_BURR_COLS", "UB_EDGE_COLS") if k in os.environ}, "ops": []}
    for name, fn, nbytes in rows:
        ms = timed(fn)
        gbs = nbytes * B / ms / 1e6
        out["ops"].append({"op": name, "ms": round(ms, 4), "us_per_frame": round(ms * 1e3 / B, 3), "algorithmic_bytes_per_frame": nbytes,
                           "achieved_gbs": round(gbs, 1), "frac_of_hbm_peak": round(gbs / peak, 4)})
        if log:
            print(f"{name:70s} {ms:8.3f} ms {ms * 1e3 / B:7.2f} us/frame {gbs:8.1f} GB/s {gbs / peak:.3f} of HBM peak", file=log)
    ctx.close()
    return out


if __name__ == "__main__":
    import torch
    torch.cuda.set_device(0)
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 512
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    print(json.dumps(measure(B, reps, log=sys.stderr)))
