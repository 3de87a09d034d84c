"""Batched two-stage inspection step and frame-sharded multi-GPU driver.

One call of :meth:`InspectionPipeline.run` is the body of the reference's hot loop
(infer_two_stage_burr.py:286-343) for a whole batch of frames: preprocess -> NestedUNet ->
argmax -> class masks -> ROI -> gray -> burr stage -> per-frame pixel counts.
Frames are independent, so across GPUs they are sharded by contiguous index ranges (one
process per GPU); the only exchange is an all-gather of the int32 [frames,4] statistics.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib
from .ops import BURR_CONFIGS, VERTICAL_ROI, map_roi_to_original


def shard_range(n_items: int, rank: int, world: int):
    """Contiguous block partition: rank r owns [r*n/world, (r+1)*n/world)."""
    return (rank * n_items) // world, ((rank + 1) * n_items) // world


def stats_line(idx, frame_no, total, cable_px, tape_px, burr_px, roi):
    """The reference's progress line (infer_two_stage_burr.py:333-343,358-359)."""
    x1, y1, x2, y2 = roi
    area = (x2 - x1) * (y2 - y1)
    cable = cable_px / area * 100 if area > 0 else 0
    tape = tape_px / area * 100 if area > 0 else 0
    burr = burr_px / area * 100 if area > 0 else 0
    status = "[BURR!]" if burr_px > 0 else "[OK]"
    return (f"[{idx:4d}] Frame {frame_no}/{total} {status} | "
            f"Cable:{cable:5.1f}% Tape:{tape:5.1f}% Burr:{burr:5.1f}%")


class InspectionPipeline:
    def __init__(self, state_dict, batch: int, frame_hw=(448, 800), net_hw=(512, 512), variant="standard",
                 sensitivity="medium", device=None, roi=None, num_classes=3, enhance=False):
        if device is None:
            device = torch.cuda.current_device()
        self.ctx = _lib.Context(device, max_batch=batch, frame_hw=frame_hw, net_hw=net_hw, num_classes=num_classes)
        self.ctx.set_weights(state_dict)
        self.device = self.ctx.device
        self.batch, self.frame_hw = batch, tuple(frame_hw)
        self.variant = _lib.VARIANT_ENHANCED if variant == "enhanced" else _lib.VARIANT_STANDARD
        if self.variant == _lib.VARIANT_ENHANCED:
            self.params = _lib.enhanced_params()
            r = VERTICAL_ROI
            self.roi = roi or (r["x1"], r["y1"], r["x2"], r["y2"])
        else:
            cfg = BURR_CONFIGS[sensitivity]
            self.params = _lib.standard_params(cfg["min_area"], cfg["max_area"])
            self.roi = roi or map_roi_to_original((frame_hw[1], frame_hw[0]), (512, 512))
        if enhance:       # --enhance of infer_enhanced_burr.py:201,279-280: enhance_image before the net and the gray conversion
            self.variant |= _lib.FLAG_ENHANCE_IMAGE

    def run(self, frames: torch.Tensor, want_masks=False):
        """frames: CUDA uint8 [B,H,W,3] BGR -> (stats int32 [B,4] on device, masks or None).
        stats columns: index in batch, cable_px, tape_px, burr_px.  Asynchronous."""
        with torch.cuda.device(self.device):
            return self.ctx.pipeline(frames, self.variant, self.roi, self.params, want_masks)

    def run_graph(self, frames: torch.Tensor, stats: torch.Tensor = None):
        """run() replayed as a CUDA graph (ub_pipeline_graph) for small batches, where the host-side launches are the step
        time.  The graph is keyed on the addresses of `frames` and `stats`: refill the same tensors between calls
        (frames.copy_(...)).  Returns the stats tensor (owned by the pipeline unless given)."""
        if stats is None:
            if getattr(self, "_graph_stats", None) is None or self._graph_stats.shape[0] != frames.shape[0]:
                self._graph_stats = torch.empty((frames.shape[0], 4), dtype=torch.int32, device=frames.device)
            stats = self._graph_stats
        with torch.cuda.device(self.device):
            self.ctx.pipeline(frames, self.variant, self.roi, self.params, stats=stats, graph=True)
        return stats

    def run_host(self, frames_host, stats_host=None, masks_host=None):
        """frames_host: (pinned) uint8 [B,H,W,3] on the host -> stats int32 [B,4] on the host.
        Includes the host->device and device->host copies; synchronous."""
        b = frames_host.shape[0]
        if stats_host is None:
            stats_host = torch.empty((b, 4), dtype=torch.int32).pin_memory()
        with torch.cuda.device(self.device):
            self.ctx.pipeline_host(frames_host, stats_host, self.variant, self.roi, self.params, masks_host)
        return stats_host

    def stream_host(self, batches, src_rotate=None):
        """Generator over an iterable of pinned host uint8 [b,H,W,3] batches -> int32 [b,4] host statistics per batch,
        in order.  Two batches are in flight: the host->device copy of batch i+1 overlaps the compute of batch i
        (ub_pipeline_submit / ub_pipeline_wait).  src_rotate = True / False: the batches are at the SOURCE resolution
        [b,sh,sw,3] and are rotated (or not) and resized to the frame size on the device (ub_pipeline_submit_src)."""
        stats = [torch.empty((self.batch, 4), dtype=torch.int32).pin_memory() for _ in range(2)]
        pending = None                                   # (slot, frames kept alive, n)
        i = 0
        with torch.cuda.device(self.device):
            for frames in batches:
                slot = i & 1
                if src_rotate is None:
                    self.ctx.pipeline_submit(frames, stats[slot], slot, self.variant, self.roi, self.params)
                else:
                    self.ctx.pipeline_submit_src(frames, stats[slot], slot, src_rotate, self.variant, self.roi, self.params)
                if pending is not None:
                    self.ctx.pipeline_wait(pending[0])
                    yield stats[pending[0]][: pending[2]].clone()
                pending = (slot, frames, frames.shape[0])
                i += 1
            if pending is not None:
                self.ctx.pipeline_wait(pending[0])
                yield stats[pending[0]][: pending[2]].clone()

    def saturated(self, reset=False) -> int:
        """Conv outputs clamped to the largest finite 16-bit value while the check was on (the first step after the weights
        were set, or always with ctx.set_option('check_saturation', 1)): non-zero means the checkpoint's activations
        overflow the operand type and the masks are not to be trusted."""
        return self.ctx.get_stat("saturated", reset)

    class AsyncGather:
        """All-gather of one step's statistics on a side stream: the collective of step i runs beside the
        compute of step i+1.  start(stats) enqueues, result() makes the caller's stream wait and returns the tensor."""

        def __init__(self, frame_offset, group=None):
            self.offset, self.group = frame_offset, group
            self.stream = torch.cuda.Stream()
            self.done = torch.cuda.Event()
            self.out = None

        def start(self, stats):
            ready = torch.cuda.Event()
            ready.record()
            with torch.cuda.stream(self.stream):
                self.stream.wait_event(ready)
                self.out = InspectionPipeline.gather_stats(stats, self.offset, self.group, equal_counts=True)
                stats.record_stream(self.stream)
                self.done.record(self.stream)

        def result(self):
            torch.cuda.current_stream().wait_event(self.done)
            return self.out

    @staticmethod
    def gather_stats(stats: torch.Tensor, frame_offset: int, group=None, equal_counts: bool = False) -> torch.Tensor:
        """All-gather per-frame statistics across ranks (NCCL on GPU tensors, gloo on CPU
        tensors).  Column 0 becomes the global frame index.  Returns [sum B_r, 4] sorted by frame.
        equal_counts=True (every rank holds the same number of frames, offsets ascending with the rank)
        skips the size exchange and the sort: one collective, no host synchronisation."""
        import torch.distributed as dist
        s = stats.clone()
        s[:, 0] += frame_offset
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return s
        world = dist.get_world_size(group)
        if equal_counts:
            out = torch.empty((world * s.shape[0], 4), dtype=s.dtype, device=s.device)
            dist.all_gather_into_tensor(out, s, group=group)
            return out
        n = torch.tensor([s.shape[0]], dtype=torch.int64, device=s.device)
        sizes = [torch.zeros_like(n) for _ in range(world)]
        dist.all_gather(sizes, n, group=group)
        mx = int(max(int(x) for x in sizes))
        pad = torch.full((mx, 4), -1, dtype=s.dtype, device=s.device)
        pad[: s.shape[0]] = s
        parts = [torch.empty_like(pad) for _ in range(world)]
        dist.all_gather(parts, pad, group=group)
        out = torch.cat([p[: int(k)] for p, k in zip(parts, sizes)], dim=0)
        return out[torch.argsort(out[:, 0])]

    def summary(self, all_stats) -> dict:
        """Counters of infer_two_stage_burr.py:321-323,369-372 from gathered stats."""
        a = all_stats.cpu().numpy() if isinstance(all_stats, torch.Tensor) else np.asarray(all_stats)
        return {"frames": int(a.shape[0]), "burr_frames": int((a[:, 3] > 0).sum()),
                "cable_px": int(a[:, 1].sum()), "tape_px": int(a[:, 2].sum()), "burr_px": int(a[:, 3].sum())}
