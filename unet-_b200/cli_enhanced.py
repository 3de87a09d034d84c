"""Drop-in for `python infer_enhanced_burr.py ...` (reference infer_enhanced_burr.py:195-366):

    python -m unet_b200.cli_enhanced --video V --model M --output DIR [--enhance]

Same flags (:197-201) and printed lines (:215-221,:224,:230,:235,:244,:250-253,:347-348,:361-365);
every frame is rotated 90 degrees counter-clockwise and resized to 800x448 (:273-276), optionally
enhanced (:279-280: CLAHE + non-local means + sharpen), segmented, and inspected with the
multi-scale detector inside the fixed ROI x in [200,600).  Differences from the reference are those
listed in cli_two_stage.py (no window, batched, no CPU fallback, --batch / --synthetic N /
--no-video-out); --synthetic generates 1080x1920-class frames so the rotate+resize path runs.
"""
from __future__ import annotations

import argparse
import sys
import time
from pathlib import Path

from . import cli_common as K
from .ops import VERTICAL_ROI

TARGET_W, TARGET_H = 800, 448


def build_parser():
    p = argparse.ArgumentParser(description="增强版毛刺检测（图像增强+改进算法）")
    p.add_argument("--video", type=str, default=None)                     # required unless --synthetic
    p.add_argument("--model", type=str, default="checkpoints_3class_advanced/best_model.pth")
    p.add_argument("--output", type=str, required=True)
    p.add_argument("--device", type=str, default="cuda")
    p.add_argument("--enhance", action="store_true", help="启用图像增强")
    p.add_argument("--batch", type=int, default=16, help="frames per device step")
    p.add_argument("--synthetic", type=int, default=0, help="use N generated portrait frames instead of --video")
    p.add_argument("--synthetic-size", type=str, default="800x448", help="HxW of the generated frames before rotation")
    p.add_argument("--no-video-out", action="store_true", help="skip the overlay drawing and result.mp4")
    return p


def main(argv=None, out=print, state_dict=None):
    p = build_parser()
    args = p.parse_args(argv)
    if args.video is None and not args.synthetic:
        p.error("the following arguments are required: --video")
    rank, world, gather = K.dist_env()
    say = out if rank == 0 else (lambda *a, **k: None)
    if world > 1 and not args.no_video_out:
        p.error("multi-process runs need --no-video-out")

    output_dir = Path(args.output)
    output_dir.mkdir(parents=True, exist_ok=True)
    R = VERTICAL_ROI
    roi = (R["x1"], R["y1"], R["x2"], R["y2"])

    say("=" * 70)
    say("增强版毛刺检测系统")
    say("=" * 70)
    say(f"图像增强: {'启用' if args.enhance else '禁用'}")
    say(f"ROI: X[{R['x1']}, {R['x2']}] Y[{R['y1']}, {R['y2']}]")
    say("毛刺检测: 多尺度边缘融合 + 形态学过滤")
    say("=" * 70)

    say(f"\n加载模型: {args.model}")
    sd = state_dict if state_dict is not None else K.load_state_dict(args.model)
    import torch
    device = K.resolve_device(args.device)
    say(f"模型已加载到 {device}")

    sh, sw = (int(v) for v in args.synthetic_size.lower().split("x"))
    src = K.open_source(args.video, args.synthetic, (sh, sw), prefetch=2 * args.batch)
    if src is None:
        say(f"错误: 无法打开视频 {args.video}")
        return 0
    total_frames, fps = src.total_frames, src.fps
    say(f"\n原始视频: {src.width}x{src.height} @ {fps:.2f}fps, 总帧数: {total_frames}")
    say(f"处理流程: 旋转90度 -> 归一化到 {TARGET_W}x{TARGET_H}")
    if args.enhance:
        say("图像增强: CLAHE + 去噪 + 锐化")
    say("-" * 70)

    writer = None
    if not args.no_video_out:
        import cv2
        writer = cv2.VideoWriter(str(output_dir / "result.mp4"), cv2.VideoWriter_fourcc(*"mp4v"), fps, (TARGET_W, TARGET_H))

    step = K.DeviceStep(sd, args.batch, (src.height, src.width), (TARGET_H, TARGET_W), rotate=True, resize=True,
                        variant="enhanced", sensitivity="medium", roi=roi, enhance=args.enhance,
                        want_masks=writer is not None, device=device.index if device.index is not None else None)
    t0 = time.time()
    processed, burr_frames, rec = K.run_loop(src, step, args.batch, roi, total_frames, 1, 60, writer,
                                             rank, world, gather, say)
    src.release()
    if writer is not None:
        writer.release()
    elapsed = time.time() - t0
    for line in K.summary_lines(processed, burr_frames, elapsed, output_dir):
        say(line)
    return rec


if __name__ == "__main__":
    main()
    sys.exit(0)
