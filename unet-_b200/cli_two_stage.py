"""Drop-in for `python infer_two_stage_burr.py ...` (reference infer_two_stage_burr.py:173-373):

    python -m unet_b200.cli_two_stage --video V --model M --output DIR [--burr-sensitivity medium] ...

Same flags and defaults (:175-187), same printed lines (:200-208,:211,:218-221,:225,:233-256,
:358-359,:369-373), same <output>/result.mp4 (:259-262).  Differences, all deliberate:
  * the imshow/waitKey window (:355,:361) is omitted (headless);
  * frames are processed --batch at a time on the device (default 64);
  * there is no CPU fallback: without a CUDA device the run fails instead of using the CPU;
  * extra optional flags: --batch, --synthetic N (generated frames instead of --video),
    --no-video-out (skip drawing and result.mp4).  Under torchrun the batches are dealt to the
    ranks round-robin and rank 0 prints; that mode needs --no-video-out.
"""
from __future__ import annotations

import argparse
import sys
import time
from pathlib import Path

from . import cli_common as K
from .ops import BURR_CONFIGS, FIXED_ROI_512, map_roi_to_original


def build_parser():
    p = argparse.ArgumentParser(description="两阶段毛刺检测")
    p.add_argument("--video", type=str, default=None)                     # required unless --synthetic
    p.add_argument("--model", type=str, default="checkpoints_3class_advanced/best_model.pth")
    p.add_argument("--output", type=str, required=True)
    p.add_argument("--device", type=str, default="cuda")
    p.add_argument("--frame-stride", type=int, default=1)
    p.add_argument("--print-interval", type=int, default=60)
    p.add_argument("--burr-sensitivity", type=str, default="medium", choices=["low", "medium", "high"])
    p.add_argument("--rotate", action="store_true", help="旋转视频90度（逆时针）")
    p.add_argument("--normalize-resolution", action="store_true",
                   help="将视频归一化到标准分辨率（800x448），适用于高分辨率视频")
    p.add_argument("--target-width", type=int, default=800, help="归一化目标宽度")
    p.add_argument("--target-height", type=int, default=448, help="归一化目标高度")
    # additions (defaults reproduce the reference's output)
    p.add_argument("--batch", type=int, default=64, help="frames per device step")
    p.add_argument("--synthetic", type=int, default=0, help="use N generated 800x448 cable frames instead of --video")
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
    cfg = BURR_CONFIGS[args.burr_sensitivity]

    say("=" * 70)
    say("两阶段毛刺检测系统")
    say("=" * 70)
    say(f"阶段1: 电缆/胶带分割 (模型: {args.model})")
    say(f"阶段2: 毛刺检测 (灵敏度: {args.burr_sensitivity})")
    say(f"  - Laplacian阈值: {cfg['laplacian_threshold']}")
    say(f"  - 检测带宽度: {cfg['band_out']}px")
    say(f"  - 面积范围: {cfg['min_area']}-{cfg['max_area']}px")
    say("=" * 70)

    say(f"\n加载模型: {args.model}")
    sd = state_dict if state_dict is not None else K.load_state_dict(args.model)
    import torch
    device = K.resolve_device(args.device)
    say(f"模型已加载到 {device}")

    r = FIXED_ROI_512
    say(f"\n固定ROI（512x512）: X[{r['x1']}, {r['x2']}] Y[{r['y1']}, {r['y2']}]")

    src = K.open_source(args.video, args.synthetic, prefetch=2 * args.batch)
    if src is None:
        say(f"错误: 无法打开视频 {args.video}")
        return 0
    total_frames, fps = src.total_frames, src.fps
    say(f"原始视频: {src.width}x{src.height}")
    w_rot, h_rot = (src.height, src.width) if args.rotate else (src.width, src.height)
    if args.rotate:
        say(f"旋转后: {w_rot}x{h_rot}")
    if args.normalize_resolution:
        width, height = args.target_width, args.target_height
        say(f"归一化到: {width}x{height}")
    else:
        width, height = w_rot, h_rot
    roi = map_roi_to_original((width, height), (512, 512))
    say(f"固定ROI（{width}x{height}）: X[{roi[0]}, {roi[2]}] Y[{roi[1]}, {roi[3]}]")
    say(f"处理分辨率: {width}x{height} @ {fps:.2f}fps, 总帧数: {total_frames}")
    say("-" * 70)

    writer = None
    if not args.no_video_out:
        import cv2
        writer = cv2.VideoWriter(str(output_dir / "result.mp4"), cv2.VideoWriter_fourcc(*"mp4v"), fps, (width, height))

    step = K.DeviceStep(sd, args.batch, (src.height, src.width), (height, width), rotate=args.rotate,
                        resize=args.normalize_resolution, variant="standard", sensitivity=args.burr_sensitivity,
                        roi=roi, enhance=False, want_masks=writer is not None,
                        device=device.index if device.index is not None else None)
    t0 = time.time()
    processed, burr_frames, rec = K.run_loop(src, step, args.batch, roi, total_frames, args.frame_stride,
                                             args.print_interval, writer, rank, world, gather, say)
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
