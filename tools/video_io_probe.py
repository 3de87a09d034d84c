"""SURVEY 8f row 2 (video ingest / egress): what this box offers and where the ceiling is.

1. Probes the NVDEC / NVENC user-mode libraries (compute-side libraries, not graphics) with ctypes.CDLL -- one line each.
2. Writes a synthetic 1080p clip with cv2.VideoWriter (mp4v, as infer_two_stage_burr.py:258-262) and measures
   cv2.VideoCapture decode frames/s (infer_two_stage_burr.py:223-231,270-272) alone: the ceiling of any loop fed by it.
3. Runs the two-stage CLI on that clip (rotate + normalise, no result video) with the decode ring and prints its frames/s
   next to the decode ceiling and the ring's own decode figure.

Usage: python tools/video_io_probe.py [frames] > profiles/rNN_video_io.json"""
import ctypes
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def probe(name):
    try:
        ctypes.CDLL(name)
        return "present"
    except OSError as e:
        return f"absent ({str(e).splitlines()[0][:80]})"


def main():
    import cv2
    import numpy as np
    import torch

    import unet_b200.synth as S
    from unet_b200 import cli_common as K
    from unet_b200 import cli_two_stage

    n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    out = {"libs": {lib: probe(lib) for lib in ("libnvcuvid.so.1", "libnvcuvid.so", "libnvidia-encode.so.1", "libnvidia-encode.so")},
           "cv2": cv2.__version__, "cv2_videoio_backends": [cv2.videoio_registry.getBackendName(b) for b in cv2.videoio_registry.getBackends()],
           "cv2_cuda_devices": cv2.cuda.getCudaEnabledDeviceCount() if hasattr(cv2, "cuda") else 0}
    tmp = tempfile.mkdtemp()
    clip = os.path.join(tmp, "clip.mp4")
    base = S.frames_cable(16, 1080, 1920, seed=3)
    t0 = time.perf_counter()
    w = cv2.VideoWriter(clip, cv2.VideoWriter_fourcc(*"mp4v"), 30.0, (1920, 1080))
    for i in range(n):
        w.write(base[i % 16])
    w.release()
    out["encode"] = {"codec": "mp4v", "frames": n, "size": "1920x1080", "frames_per_s": n / (time.perf_counter() - t0),
                     "bytes": os.path.getsize(clip)}
    cap = cv2.VideoCapture(clip)
    t0 = time.perf_counter()
    k = 0
    while True:
        ok, _ = cap.read()
        if not ok:
            break
        k += 1
    out["decode_alone"] = {"frames": k, "frames_per_s": k / (time.perf_counter() - t0), "threads": cv2.getNumThreads()}
    cap.release()
    if torch.cuda.is_available():
        sd = S.make_state_dict(0)
        for ring in (None, True, False):            # None: warm-up run (context creation, lazy tables), not reported
            orig = K.open_source
            holder = {}

            def opener(video, synthetic, *a, prefetch=0, **kw):
                holder["src"] = orig(video, synthetic, *a, prefetch=prefetch if ring is not False else 0, **kw)
                return holder["src"]
            K.open_source = opener
            try:
                lines = []
                t0 = time.perf_counter()
                cli_two_stage.main(["--video", clip, "--model", "unused", "--output", tmp, "--rotate", "--normalize-resolution",
                                    "--no-video-out", "--print-interval", "100000"], state_dict=sd, out=lines.append)
                dt = time.perf_counter() - t0
            finally:
                K.open_source = orig
            if ring is None:
                continue
            row = {"frames_per_s_wall": k / dt, "summary": [l for l in lines if "FPS" in l][-1:]}
            if ring:
                row["ring_decode_frames_per_s"] = holder["src"].decode_fps()
            out["cli_two_stage_1080p_rotate_normalize" + ("_decode_ring" if ring else "_no_ring")] = row
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
