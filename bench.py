#!/usr/bin/env python
"""Benchmark of the per-frame inspection hot path (BASELINE.json metric: frames/sec,
800x448 UNet++ + burr).

  python bench.py [--gpus N] [--steps K] [--warmup W]          # this repo, one rank per GPU
  python bench.py --impl reference ...                         # reference CPU path on host cores

A "step" is one pass of the whole two-stage pipeline (preprocess -> NestedUNet 512x512 ->
argmax/masks/ROI -> gray -> burr stage -> per-frame counts) over one batch of synthetic
800x448 frames.  `value` is measured with the frames already resident in HBM; `e2e` is the
same steps through the C-ABI host entry points (ub_pipeline_submit / ub_pipeline_wait, two batches
in flight) with pinned HOST frames in and HOST statistics out, every step's copies inside the
timed region.  One JSON line is printed by rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FRAME_HW = (448, 800)
NET_HW = (512, 512)
GFLOP_PER_FRAME = 111.636          # 18 conv3x3 + final at 512x512, 2*MACs
METRIC = "frames/sec (800x448 UNet++ + burr)"
CONV_DRAM_BYTES_PER_FRAME = (11.661369e9 + 7.043801e9) / 64     # measured, see roofline.traffic


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "tf": d["bf16_tflops"], "tf_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                "source": "MEASURED_PEAKS.json"}
    return {"hbm_gbs": 6650.0, "tf": 1590.0, "tf_sustained": 1400.0, "source": "fallback (B200_PROFILING.md)"}


class ClockSampler:
    """Samples SM clock and clock-event (throttle) reasons during the timed region: NVML polled every 20 ms from a
    thread (the timed region is ~0.5 s, shorter than nvidia-smi's start-up), nvidia-smi -lms as the fallback."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    BITS = (("sw_power_cap", 0x4), ("hw_slowdown", 0x8), ("sw_thermal_slowdown", 0x20), ("hw_thermal_slowdown", 0x40))

    def __init__(self, index):
        self.index, self.rows, self.proc, self.nvml, self.stop_flag = index, [], None, None, False
        self.samples, self.mask, self.max_mhz = [], 0, None
        try:
            import pynvml
            import torch
            pynvml.nvmlInit()
            try:
                uuid = str(torch.cuda.get_device_properties(index).uuid)
                self.h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
            except Exception:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
        except Exception:
            self.nvml = None

    def _poll(self):
        n = self.nvml
        reasons = getattr(n, "nvmlDeviceGetCurrentClocksEventReasons", None) or n.nvmlDeviceGetCurrentClocksThrottleReasons
        while not self.stop_flag:
            try:
                self.samples.append(float(n.nvmlDeviceGetClockInfo(self.h, n.NVML_CLOCK_SM)))
                self.mask |= int(reasons(self.h))
            except Exception:
                pass
            time.sleep(0.02)

    def start(self):
        if self.nvml is not None:
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.nvml is not None:
            self.stop_flag = True
            self.t.join(timeout=2)
            sm = sorted(self.samples)
            return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.max_mhz,
                    "reasons": sorted(name for name, bit in self.BITS if self.mask & bit), "samples": len(sm), "source": "nvml"}
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm), "source": "nvidia-smi"}


def make_frames(batch, seed, hw=FRAME_HW):
    import numpy as np
    import unet_b200.synth as S
    # two thirds structured cable frames, one third speckle (non-trivial burr output); generated 64 at a time and tiled
    base = min(batch, 64)
    n_sp = base // 3
    parts = [S.frames_cable(base - n_sp, *hw, seed=seed)]
    if n_sp:
        parts.append(S.frames_speckle(n_sp, *hw, seed=seed + 1000))
    one = np.concatenate(parts)
    return np.concatenate([one] * ((batch + base - 1) // base))[:batch]


def workload_hw(workload):
    return (512, 512) if workload == "net512" else FRAME_HW


def ref_step(R, frame, sd, workload):
    """One iteration of the reference loop body for the workload (oracle/ref_pipeline.py)."""
    if workload == "enhanced":        # BASELINE config 4: rotate + resize + enhance_image + multi-scale detector
        fr = R.normalize_frame(frame, rotate=True, size=(FRAME_HW[1], FRAME_HW[0]))
        return R.frame_step(fr, sd, variant="enhanced", do_enhance=True)
    return R.frame_step(frame, sd, "medium")


def make_workload_frames(workload, batch, seed):
    if workload == "enhanced":
        import unet_b200.synth as S
        return S.frames_cable(batch, 1080, 1920, seed=seed)        # SURVEY 8(d) `hires`
    return make_frames(batch, seed, workload_hw(workload))


def cpu_reference_fps(frames, sd, n_frames, warm=1, workload="two_stage"):
    """Times the oracle restatement of the reference loop body (torch CPU + OpenCV, the same
    library calls the reference makes) on the host cores.  Checker-side code: only the
    cpu_baseline / --impl reference legs run it."""
    import cv2
    import torch
    from oracle import ref_pipeline as R
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cv2.setNumThreads(cores)
    for i in range(warm):
        ref_step(R, frames[i % len(frames)], sd, workload)
    t0 = time.perf_counter()
    for i in range(n_frames):
        ref_step(R, frames[i % len(frames)], sd, workload)
    dt = time.perf_counter() - t0
    return n_frames / dt, cores, dt


def run_reference(args):
    """--impl reference: the reference's CPU path for the same metric, on the host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import unet_b200.synth as S
    sd = S.make_state_dict(0)
    per_step = args.ref_frames_per_step
    frames = make_workload_frames(args.workload, max(per_step, 3), seed=0)
    import cv2
    import torch
    from oracle import ref_pipeline as R
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cv2.setNumThreads(cores)
    k = 0
    for _ in range(args.warmup):
        for _ in range(per_step):
            ref_step(R, frames[k % len(frames)], sd, args.workload); k += 1
    t0 = time.perf_counter()
    for _ in range(args.steps):
        for _ in range(per_step):
            ref_step(R, frames[k % len(frames)], sd, args.workload); k += 1
    dt = time.perf_counter() - t0
    fps = args.steps * per_step / dt
    sample = f"{per_step} frames/step of the {args.batch}-frame batch, batch 1, fp32, torch {torch.__version__} + cv2 {cv2.__version__}"
    line = {"impl": "reference", "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(args.workload, args.batch),
                       "frame": f"{workload_hw(args.workload)[1]}x{workload_hw(args.workload)[0]}", "net": "512x512",
                       "batch_per_gpu": args.batch},
            "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def workload_name(workload, batch):
    if workload == "enhanced":
        return f"enhanced_pipeline_1080p_rotate_resize_enhance_800x448_batch{batch}"
    if workload == "net512":          # BASELINE config 5: 512x512 frames, batch 256, statistics gathered over NCCL
        return f"two_stage_pipeline_512x512_batch{batch}"
    return f"two_stage_pipeline_800x448_batch{batch}"


def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import unet_b200.synth as S
    from unet_b200.pipeline import InspectionPipeline, shard_range

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    B = args.batch
    sd = S.make_state_dict(0)
    enhanced = args.workload == "enhanced"
    fhw = workload_hw(args.workload)
    pipe = InspectionPipeline(sd, batch=B, frame_hw=fhw, net_hw=NET_HW, device=local,
                              variant="enhanced" if enhanced else "standard", enhance=enhanced)
    # every rank owns a contiguous block of the global frame index space (weak scaling)
    lo, _ = shard_range(B * world, rank, world)
    frames_np = make_workload_frames(args.workload, B, seed=rank)
    frames_host = torch.from_numpy(frames_np).pin_memory()
    frames_dev = frames_host.cuda()
    stats_host = torch.empty((B, 4), dtype=torch.int32).pin_memory()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_dev(src=None):
        f = frames_dev if src is None else src
        if enhanced:      # 1080p source: rotate 90 degrees CCW + resize to 800x448 (infer_enhanced_burr.py:273-276)
            f = pipe.ctx.normalize_frames(f, (FRAME_HW[1], FRAME_HW[0]), rotate=True)
        stats, _ = pipe.run(f)
        if world > 1:
            return InspectionPipeline.gather_stats(stats, lo, equal_counts=True)
        return stats

    def step_host():
        if enhanced:      # host frames in, host statistics out; copies on the compute stream inside the timed region
            out = step_dev(frames_host.cuda(non_blocking=True))
            stats_host.copy_(out[:B] if world == 1 else out[lo:lo + B])
            return
        pipe.run_host(frames_host, stats_host)
        if world > 1:
            InspectionPipeline.gather_stats(stats_host.cuda(non_blocking=True), lo, equal_counts=True)

    for _ in range(args.warmup):
        out = step_dev()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    pipe.ctx.timing_enable(True)
    pipe.ctx.timing_read(reset=True)
    launches0 = pipe.ctx.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    for _ in range(args.steps):
        out = step_dev()
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    launches = pipe.ctx.launch_count() - launches0
    tm = pipe.ctx.timing_read(reset=True)
    pipe.ctx.timing_enable(False)
    clocks = sampler.stop() if rank == 0 else None
    summary = pipe.summary(out)

    # end to end through the host entry point (H2D + compute + D2H inside the timed region)
    def run_host_steps(n):
        if enhanced:
            for _ in range(n):
                step_host()
            return
        # the streaming host entry points (ub_pipeline_submit / _wait): every step copies its own frames host->device
        # and its statistics device->host; the copy of step i+1 overlaps the compute of step i
        for st_h in pipe.stream_host(frames_host for _ in range(n)):
            if world > 1:
                InspectionPipeline.gather_stats(st_h.cuda(non_blocking=True), lo, equal_counts=True)

    run_host_steps(max(2, args.warmup // 2))
    barrier()
    t0 = time.perf_counter()
    run_host_steps(args.steps)
    barrier()
    e2e_s = time.perf_counter() - t0

    t = torch.tensor([ms, e2e_s * 1e3], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_ms = float(t[0]), float(t[1])
    if rank == 0:
        peaks = load_peaks()
        fps = B * world * args.steps / (ms / 1e3)
        conv_ms, conv_n = tm["conv3x3"]
        tflops = GFLOP_PER_FRAME * B * args.steps / conv_ms if conv_ms > 0 else 0.0     # GFLOP/ms = TFLOP/s
        roof = {"bound": "tensor", "kernel": "k_conv_tc (18 launches/step, tcgen05 implicit GEMM)",
                "achieved": tflops, "peak": peaks["tf_sustained"], "unit": "TFLOP/s",
                "frac": tflops / peaks["tf_sustained"], "peak_source": peaks["source"] + " bf16_tflops_sustained",
                "frac_of_burst": tflops / peaks["tf"],
                # dram__bytes_read.sum + dram__bytes_write.sum over the 18 conv launches of one 64-frame step
                # (ncu --set full, profiles/r01i_conv_ncu_full.csv): 11.66 GB read + 7.04 GB written
                "traffic": CONV_DRAM_BYTES_PER_FRAME * B, "traffic_unit": "bytes per step (18 launches), from profiles/r01i_conv_ncu_full.csv",
                "algorithmic": f"{GFLOP_PER_FRAME} GFLOP/frame x {B} frames per step over the 18 conv launches",
                "ms_per_step_by_class": {k: v[0] / args.steps for k, v in tm.items()},
                "ms_per_step_by_class_note": "CUDA-event brackets per kernel class; gray/edge-map classes run on a second stream "
                                             "beside the convs, so the classes overlap and do not add up to ms_per_step"}
        line = {"metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f16", "data": "synthetic",
                "config": {"workload": workload_name(args.workload, B), "frame": f"{fhw[1]}x{fhw[0]}", "net": "512x512",
                           "batch_per_gpu": B, "parallelism": f"frames sharded over {world} GPU(s)",
                           "l2": "activation working set ~12 GB/step per GPU >> 126 MB L2; no explicit flush"},
                "e2e": {"value": B * world * args.steps / (e2e_ms / 1e3), "unit": "frames/s",
                        "h2d_bytes_per_step": int(frames_host.numel()), "d2h_bytes_per_step": int(stats_host.numel() * 4)},
                "gpu_launches": int(launches), "roofline": roof, "clocks": clocks,
                "checks": {"burr_frames": summary["burr_frames"], "frames": summary["frames"]}}
        if world == 1 and not args.no_cpu_baseline:
            fps_cpu, cores, dt = cpu_reference_fps(frames_np, sd, args.cpu_frames, workload=args.workload)
            line["cpu_baseline"] = {"value": fps_cpu, "unit": "frames/s", "cores": cores, "kind": "port",
                                    "sample": f"{args.cpu_frames} frames of the same batch, batch 1, fp32 ({dt:.1f} s)"}
        if world == 1 and not args.no_hbm_kernels and args.workload == "two_stage":
            # the HBM-bound side of the path, one entry point at a time on 512 resident frames (planes > L2)
            torch.cuda.empty_cache()
            sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "tools"))
            import bench_stage2
            line["hbm_kernels"] = bench_stage2.measure(512, 10)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=64, help="frames per step per GPU")
    ap.add_argument("--cpu-frames", type=int, default=24)
    ap.add_argument("--ref-frames-per-step", type=int, default=2)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-hbm-kernels", action="store_true", help="skip the per-entry-point stage-2 HBM table")
    ap.add_argument("--workload", default="two_stage", choices=["two_stage", "enhanced", "net512"],
                    help="two_stage: BASELINE metric config (default); enhanced: config 4 (1080p rotate+resize, enhance_image, "
                         "multi-scale detector); net512: config 5 (512x512 frames, batch 256 per GPU)")
    args = ap.parse_args()
    if args.workload == "enhanced" and args.batch == 64:
        args.batch = 16
    if args.workload == "net512" and args.batch == 64:
        args.batch = 256
    # The contract is ONE JSON line on stdout.  Native libraries write there too (NCCL prints its version banner
    # from C), so fd 1 is pointed at stderr for the run and the JSON line goes to the saved descriptor.
    sys.stdout.flush()
    real_out = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    global print
    _print = print
    print = lambda *a, **k: _print(*a, **{**k, "file": real_out})      # noqa: A001  (both JSON prints below)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
