"""Shared machinery of the two drop-in command lines (cli_two_stage.py, cli_enhanced.py).

The reference scripts (infer_two_stage_burr.py:173-373, infer_enhanced_burr.py:195-366) read one
frame, run it, draw it, print every N-th line.  Here the same loop is batched: frames are
collected `--batch` at a time into pinned host memory, normalised / enhanced / segmented /
inspected on the device in one go, and the per-frame statistics come back as int32 [B,4].
Video decode/encode and the overlay drawing stay on the host with cv2 (they are I/O, not the
hot path) and are skipped with --no-video-out.  Under torchrun every rank takes every
world-th batch; the statistics are all-gathered and rank 0 prints the reference's lines.
"""
from __future__ import annotations

import os
import time

import numpy as np

CLASS_COLORS = {0: (0, 0, 0), 1: (0, 255, 0), 2: (255, 0, 0), 3: (255, 0, 255)}   # BGR, as the reference's table


# ------------------------------------------------------------------------------------- sources
class ArraySource:
    """Frames held in memory ([N,H,W,3] uint8): --synthetic N or a .npy file."""

    def __init__(self, frames: np.ndarray, fps: float = 30.0):
        assert frames.ndim == 4 and frames.shape[3] == 3 and frames.dtype == np.uint8
        self.frames, self.fps, self.pos = frames, fps, 0
        self.total_frames = int(frames.shape[0])
        self.height, self.width = int(frames.shape[1]), int(frames.shape[2])

    def read(self):
        if self.pos >= self.total_frames:
            return None
        f = self.frames[self.pos]
        self.pos += 1
        return f

    def release(self):
        pass


class VideoSource:
    """cv2.VideoCapture, as infer_two_stage_burr.py:222-231."""

    def __init__(self, path: str):
        import cv2
        self.cap = cv2.VideoCapture(path)
        self.ok = self.cap.isOpened()
        if self.ok:
            self.total_frames = int(self.cap.get(cv2.CAP_PROP_FRAME_COUNT))
            self.fps = self.cap.get(cv2.CAP_PROP_FPS)
            self.width = int(self.cap.get(cv2.CAP_PROP_FRAME_WIDTH))
            self.height = int(self.cap.get(cv2.CAP_PROP_FRAME_HEIGHT))

    def read(self):
        ret, frame = self.cap.read()
        return frame if ret else None

    def release(self):
        self.cap.release()


class PrefetchSource:
    """Decode ring: a thread reads ahead of the loop into a bounded queue, so that cv2's decode (which releases the GIL)
    overlaps the device step, the drawing and the encode of the previous batch instead of alternating with them as in the
    reference loop (infer_two_stage_burr.py:268-272).  Same read() / release() contract and frame order as the source."""

    def __init__(self, src, depth: int = 128):
        import queue
        import threading
        self.src = src
        self.total_frames, self.fps, self.width, self.height = src.total_frames, src.fps, src.width, src.height
        self.q = queue.Queue(maxsize=max(2, depth))
        self.stop = False
        self.done = False
        self.decode_s = 0.0                       # time the thread spent inside src.read(): the decode ceiling
        self.frames_read = 0
        self.t = threading.Thread(target=self._run, daemon=True)
        self.t.start()

    def _run(self):
        while not self.stop:
            t0 = time.perf_counter()
            f = self.src.read()
            self.decode_s += time.perf_counter() - t0
            if f is not None:
                self.frames_read += 1
            while not self.stop:
                try:
                    self.q.put(f, timeout=0.1)
                    break
                except Exception:
                    continue
            if f is None:
                return

    def read(self):
        if self.done:
            return None
        f = self.q.get()
        if f is None:
            self.done = True
        return f

    def decode_fps(self):
        return self.frames_read / self.decode_s if self.decode_s > 0 else float("inf")

    def release(self):
        self.stop = True
        try:
            while True:
                self.q.get_nowait()
        except Exception:
            pass
        self.t.join(timeout=5)
        self.src.release()


def open_source(video, synthetic: int, synthetic_hw=(448, 800), seed: int = 0, prefetch: int = 0):
    """-> source or None (None = cannot be opened; the caller prints the reference's error line).  prefetch > 0 wraps a
    video file in a decode ring of that many frames."""
    src = _open_source(video, synthetic, synthetic_hw, seed)
    if src is not None and prefetch > 0 and isinstance(src, VideoSource):
        return PrefetchSource(src, prefetch)
    return src


def _open_source(video, synthetic: int, synthetic_hw=(448, 800), seed: int = 0):
    if synthetic:
        from . import synth
        return ArraySource(synth.frames_cable(synthetic, synthetic_hw[0], synthetic_hw[1], seed=seed))
    if video is None:
        return None
    if video.endswith(".npy"):
        if not os.path.exists(video):
            return None
        return ArraySource(np.load(video, mmap_mode="r"))
    src = VideoSource(video)
    return src if src.ok else None


def load_state_dict(path):
    """torch.load(path)['model'] as infer_two_stage_burr.py:215-216 (a missing file raises).  Also accepts the
    other checkpoint layouts the reference's scripts read (infer_video.py:148-153): 'model_state_dict', or a bare
    state_dict."""
    import torch
    checkpoint = torch.load(path, map_location="cpu", weights_only=False)
    if isinstance(checkpoint, dict):
        for key in ("model", "model_state_dict"):
            if key in checkpoint and isinstance(checkpoint[key], dict):
                return checkpoint[key]
    return checkpoint


# ------------------------------------------------------------------------------------- drawing
def draw_overlay(blended, mask_cable, mask_tape, mask_burr, roi, draw_roi_box=True, redo=None):
    """Second half of the reference's visualisation (infer_two_stage_burr.py:156-170, infer_enhanced_burr.py:177-190) over the
    frame the device already blended (ub_overlay).  redo: None -- nothing annotated on the device, cv2 draws box and contours;
    0 -- the device painted everything (ops.annotate_on_device); 1 -- it did, but cv2 must redraw the burr contours on top."""
    from .ops import draw_host_contours
    img = np.ascontiguousarray(blended)
    if redo is None:
        return draw_host_contours(img, mask_burr, mask_cable, mask_tape, roi if draw_roi_box else None)
    return draw_host_contours(img, mask_burr) if redo else img


def draw_text(img, info_text, defect_text):
    import cv2
    for text, y in ((info_text, 30), (defect_text, 60)):
        cv2.putText(img, text, (10, y), cv2.FONT_HERSHEY_SIMPLEX, 0.7, (255, 255, 255), 2)
        cv2.putText(img, text, (10, y), cv2.FONT_HERSHEY_SIMPLEX, 0.7, (0, 0, 0), 1)


def ratios(cable_px, tape_px, burr_px, roi):
    x1, y1, x2, y2 = roi
    area = (x2 - x1) * (y2 - y1)
    f = (lambda v: v / area * 100) if area > 0 else (lambda v: 0)
    return f(cable_px), f(tape_px), f(burr_px)


def progress_line(processed, frame_no, total, cable_px, tape_px, burr_px, roi):
    """infer_two_stage_burr.py:358-359 / infer_enhanced_burr.py:347-348."""
    c, t, b = ratios(cable_px, tape_px, burr_px, roi)
    status = "[BURR!]" if burr_px > 0 else "[OK]"
    return (f"[{processed:4d}] Frame {frame_no}/{total} {status} | "
            f"Cable:{c:5.1f}% Tape:{t:5.1f}% Burr:{b:5.1f}%")


def summary_lines(processed, burr_frames, elapsed, output_dir):
    """infer_two_stage_burr.py:369-373."""
    return ["-" * 70,
            f"完成! 处理了 {processed} 帧, 用时 {elapsed:.1f}s, 平均 FPS: {processed/elapsed:.2f}",
            f"检测到毛刺的帧数: {burr_frames}/{processed} ({burr_frames/processed*100:.1f}%)",
            f"结果已保存到: {output_dir}"]


# ------------------------------------------------------------------------------------- the loop
class DeviceStep:
    """The per-batch device work: (rotate, resize) -> (enhance) -> two-stage pipeline."""

    def __init__(self, state_dict, batch, src_hw, proc_hw, rotate, resize, variant, sensitivity, roi, enhance,
                 want_masks, device=None):
        import torch
        from . import _lib
        from .pipeline import InspectionPipeline
        if not torch.cuda.is_available():
            raise _lib.UbError("no CUDA device: unet_b200 has no CPU fallback (the reference would fall back to the CPU here)")
        self.torch = torch
        self.pipe = InspectionPipeline(state_dict, batch=batch, frame_hw=proc_hw, variant=variant, sensitivity=sensitivity,
                                       device=device, roi=roi)
        self.rotate, self.resize, self.enhance, self.want_masks = rotate, resize, enhance, want_masks
        self.proc_hw = tuple(proc_hw)
        self.dev = self.pipe.device
        self.h_in = torch.empty((batch,) + tuple(src_hw) + (3,), dtype=torch.uint8).pin_memory()
        self.h_np = self.h_in.numpy()

    def __call__(self, frames_np):
        """frames_np: list of [h,w,3] uint8 host frames (source resolution).
        -> stats int32 [n,4] (numpy), and if want_masks (blended frames [n,H,W,3], masks [n,3,H,W]) on the host."""
        torch = self.torch
        n = len(frames_np)
        for i, f in enumerate(frames_np):
            self.h_np[i] = f
        with torch.cuda.device(self.dev):
            d = self.h_in[:n].to(self.dev, non_blocking=True)
            if self.rotate or self.resize:
                d = self.pipe.ctx.normalize_frames(d, (self.proc_hw[1], self.proc_hw[0]), self.rotate)
            if self.enhance:
                d = self.pipe.ctx.enhance(d)
            stats, masks = self.pipe.run(d, want_masks=self.want_masks)
            stats = stats.cpu().numpy()
            if self.want_masks:          # blended frames for the result video: the addWeighted chain runs on the device
                from .ops import annotate_on_device
                mc, mt = masks[:, 0].contiguous(), masks[:, 1].contiguous()
                blend = self.pipe.ctx.overlay(d, mc, mt, masks[:, 2].contiguous(), self.pipe.roi)
                # ROI box + label and the three contour sets on the device too (run_loop redraws flagged frames' burr contours)
                redo = annotate_on_device(self.pipe.ctx, blend, mc, mt, masks[:, 2].contiguous(), self.pipe.roi)
                self.redo = redo.cpu().numpy() if redo is not None else None
                return stats, blend.cpu().numpy(), masks.cpu().numpy()
        return stats, None, None


def run_loop(source, step, batch, roi, total_frames, frame_stride=1, print_interval=60, writer=None,
             rank=0, world=1, gather=None, out=print):
    """Drives `step` over the source.  `step(list_of_frames) -> (stats[n,4], frames|None, masks|None)`.
    Returns (processed_count, burr_detected_frames, records) where records is the gathered
    int64 [processed, 5] array (processed index, frame number, cable_px, tape_px, burr_px)."""
    frame_count = processed_count = 0
    pending, meta, records = [], [], []
    batch_index = 0
    t0 = time.time()

    def flush():
        nonlocal pending, meta, batch_index
        if not pending:
            return
        mine = batch_index % world == rank
        batch_index += 1
        if mine:
            stats, frames, masks = step(pending)
            for i, (proc_no, frame_no) in enumerate(meta):
                cable_px, tape_px, burr_px = (int(v) for v in stats[i, 1:4])
                records.append((proc_no, frame_no, cable_px, tape_px, burr_px))
                if writer is not None:
                    redo = getattr(step, "redo", None)
                    img = draw_overlay(frames[i], masks[i, 0], masks[i, 1], masks[i, 2], roi, redo=None if redo is None else int(redo[i]))
                    elapsed = time.time() - t0
                    fps_now = proc_no / elapsed if elapsed > 0 else 0
                    c, t, b = ratios(cable_px, tape_px, burr_px, roi)
                    status = "[BURR!]" if burr_px > 0 else "[OK]"
                    draw_text(img, f"Frame: {frame_no}/{total_frames} | FPS: {fps_now:.1f}",
                              f"{status} Cable:{c:.1f}% Tape:{t:.1f}% Burr:{b:.1f}%")
                    writer.write(img)
                if world == 1 and proc_no % print_interval == 0:
                    out(progress_line(proc_no, frame_no, total_frames, cable_px, tape_px, burr_px, roi))
        pending, meta = [], []

    while True:
        frame = source.read()
        if frame is None:
            break
        frame_count += 1
        if frame_count % frame_stride != 0:
            continue
        processed_count += 1
        pending.append(frame)
        meta.append((processed_count, frame_count))
        if len(pending) == batch:
            flush()
    flush()

    rec = np.array(records, np.int64).reshape(-1, 5)
    if world > 1:
        rec = gather(rec)
        rec = rec[np.argsort(rec[:, 0], kind="stable")]
        if rank == 0:
            for proc_no, frame_no, c, t, b in rec:
                if proc_no % print_interval == 0:
                    out(progress_line(int(proc_no), int(frame_no), total_frames, int(c), int(t), int(b), roi))
    burr_frames = int((rec[:, 4] > 0).sum())
    return processed_count, burr_frames, rec


def dist_env():
    """(rank, world, gather) from torchrun's environment; gather all-gathers ragged int64 [n,5] arrays."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    if world == 1:
        return 0, 1, None
    import torch
    import torch.distributed as dist
    if not dist.is_initialized():
        backend = "nccl" if torch.cuda.is_available() else "gloo"
        if torch.cuda.is_available():
            torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
        dist.init_process_group(backend)

    def gather(rec):
        dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
        n = torch.tensor([rec.shape[0]], dtype=torch.int64, device=dev)
        sizes = [torch.zeros_like(n) for _ in range(world)]
        dist.all_gather(sizes, n)
        mx = max(int(s) for s in sizes)
        pad = torch.full((mx, 5), -1, dtype=torch.int64, device=dev)
        pad[: rec.shape[0]] = torch.from_numpy(rec).to(dev)
        parts = [torch.empty_like(pad) for _ in range(world)]
        dist.all_gather(parts, pad)
        return torch.cat([p[: int(k)] for p, k in zip(parts, sizes)]).cpu().numpy()

    import atexit
    atexit.register(lambda: dist.is_initialized() and dist.destroy_process_group())
    return rank, world, gather


def resolve_device(name):
    """--device of the reference CLIs (infer_two_stage_burr.py:184,206): only CUDA devices exist here.  'cuda' resolves to
    the current device so that the printed device is the one the run uses; anything else fails loudly (no CPU path)."""
    import torch
    from . import _lib
    dev = torch.device(name)
    if dev.type != "cuda":
        raise _lib.UbError(f"--device {name}: unet_b200 has no CPU path; use cuda or cuda:N")
    if not torch.cuda.is_available():
        raise _lib.UbError("no CUDA device: unet_b200 has no CPU fallback")
    return torch.device("cuda", dev.index if dev.index is not None else torch.cuda.current_device())
