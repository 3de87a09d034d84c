"""CPU-only: the host logic of the drop-in command lines -- flag surface of the reference
(infer_two_stage_burr.py:175-187, infer_enhanced_burr.py:197-201), the batched loop's frame
accounting (stride, print interval, ragged last batch) against a per-frame restatement of the
reference loop, and the world_size-2 round-robin sharding over gloo.  The device step is replaced
by a CPU stand-in; nothing here touches the GPU."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from unet_b200 import cli_common as K
from unet_b200 import cli_enhanced, cli_two_stage

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _flags(parser):
    return {a.option_strings[0]: (a.default, a.required, tuple(a.choices) if a.choices else None)
            for a in parser._actions if a.option_strings and a.option_strings[0] != "-h"}


def test_two_stage_flags_match_reference():
    f = _flags(cli_two_stage.build_parser())
    ref = {"--model": "checkpoints_3class_advanced/best_model.pth", "--device": "cuda", "--frame-stride": 1,
           "--print-interval": 60, "--burr-sensitivity": "medium", "--rotate": False, "--normalize-resolution": False,
           "--target-width": 800, "--target-height": 448}
    for k, v in ref.items():
        assert f[k][0] == v, k
    assert f["--output"][1] is True and f["--burr-sensitivity"][2] == ("low", "medium", "high")
    assert set(f) - set(ref) - {"--video", "--output"} == {"--batch", "--synthetic", "--no-video-out"}
    with pytest.raises(SystemExit):                      # --video is required unless --synthetic
        cli_two_stage.main(["--output", "/tmp/x"])


def test_enhanced_flags_match_reference():
    f = _flags(cli_enhanced.build_parser())
    assert f["--model"][0] == "checkpoints_3class_advanced/best_model.pth" and f["--device"][0] == "cuda"
    assert f["--enhance"][0] is False and f["--output"][1] is True
    assert set(f) == {"--video", "--model", "--output", "--device", "--enhance", "--batch", "--synthetic",
                      "--synthetic-size", "--no-video-out"}


def test_progress_and_summary_lines():
    roi = (218, 0, 421, 448)
    assert K.progress_line(60, 120, 500, 9094, 4547, 0, roi) == \
        "[  60] Frame 120/500 [OK] | Cable: 10.0% Tape:  5.0% Burr:  0.0%"
    assert K.progress_line(1200, 1200, 1200, 0, 0, 91, roi).startswith("[1200] Frame 1200/1200 [BURR!] | Cable:  0.0%")
    s = K.summary_lines(10, 3, 2.0, "out")
    assert s[0] == "-" * 70 and "10" in s[1] and "5.00" in s[1] and s[2].endswith("3/10 (30.0%)") and s[3].endswith("out")


def _fake_stats(frame):
    v = int(frame.astype(np.int64).sum())
    return v % 5000, v % 777, (v % 3 == 0) * (v % 91)


def _fake_step(frames):
    st = np.array([(i,) + _fake_stats(f) for i, f in enumerate(frames)], np.int32)
    return st, None, None


def _reference_loop(frames, stride, interval, roi, total):
    """Per-frame restatement of the reference's accounting (infer_two_stage_burr.py:268-359)."""
    lines, frame_count, processed, burr_frames = [], 0, 0, 0
    for f in frames:
        frame_count += 1
        if frame_count % stride != 0:
            continue
        processed += 1
        c, t, b = _fake_stats(f)
        burr_frames += b > 0
        if processed % interval == 0:
            lines.append(K.progress_line(processed, frame_count, total, c, t, b, roi))
    return processed, burr_frames, lines


@pytest.mark.parametrize("n,batch,stride,interval", [(50, 8, 1, 7), (33, 64, 1, 10), (41, 4, 3, 2), (5, 2, 7, 1)])
def test_run_loop_matches_per_frame_loop(n, batch, stride, interval):
    rng = np.random.default_rng(n)
    frames = rng.integers(0, 256, (n, 6, 8, 3), dtype=np.uint8)
    roi = (1, 0, 5, 6)
    got = []
    processed, burr_frames, rec = K.run_loop(K.ArraySource(frames), _fake_step, batch, roi, n, stride, interval,
                                             out=got.append)
    want = _reference_loop(frames, stride, interval, roi, n)
    assert (processed, burr_frames, got) == want
    assert rec.shape == (processed, 5) and (rec[:, 0] == np.arange(1, processed + 1)).all()


def test_open_source_and_unopenable_video(tmp_path):
    assert K.open_source(str(tmp_path / "missing.mp4"), 0) is None          # reference prints an error and returns
    assert K.open_source(str(tmp_path / "missing.npy"), 0) is None
    fr = np.zeros((3, 4, 6, 3), np.uint8)
    np.save(tmp_path / "v.npy", fr)
    s = K.open_source(str(tmp_path / "v.npy"), 0)
    assert (s.total_frames, s.height, s.width) == (3, 4, 6)
    s = K.open_source(None, 2, (16, 24))
    assert (s.total_frames, s.height, s.width) == (2, 16, 24) and s.read().shape == (16, 24, 3)


def test_checkpoint_layouts(tmp_path):
    """'model' (infer_two_stage_burr.py:216), 'model_state_dict' and bare state_dicts (infer_video.py:148-153)."""
    import torch
    sd = {"final.weight": torch.ones(3, 32, 1, 1), "final.bias": torch.zeros(3)}
    for name, obj in (("a.pth", {"model": sd, "epoch": 3}), ("b.pth", {"model_state_dict": sd}), ("c.pth", sd)):
        torch.save(obj, tmp_path / name)
        got = K.load_state_dict(str(tmp_path / name))
        assert set(got) == set(sd) and torch.equal(got["final.weight"], sd["final.weight"])
    with pytest.raises(FileNotFoundError):
        K.load_state_dict(str(tmp_path / "missing.pth"))


_WORKER = r"""
import os, sys
import numpy as np
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, os.path.join(sys.argv[1], "tests"))
import torch.distributed as dist
dist.init_process_group("gloo")
from unet_b200 import cli_common as K
import test_cli_host as T
rank, world, gather = K.dist_env()
frames = np.random.default_rng(1).integers(0, 256, (37, 6, 8, 3), dtype=np.uint8)
roi = (1, 0, 5, 6)
got = []
processed, burr_frames, rec = K.run_loop(K.ArraySource(frames), T._fake_step, 4, roi, 37, 1, 5, None, rank, world, gather,
                                         got.append if rank == 0 else (lambda s: None))
want = T._reference_loop(frames, 1, 5, roi, 37)
assert (processed, burr_frames) == want[:2], (processed, burr_frames, want[:2])
if rank == 0:
    assert got == want[2], got
assert rec.shape == (37, 5)
dist.destroy_process_group()
print("ok", rank)
"""


def test_run_loop_gloo_world2(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(_WORKER)
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, str(script), ROOT], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT, text=True))
    for p in procs:
        out, _ = p.communicate(timeout=180)
        assert p.returncode == 0, out


def test_device_flag_has_no_cpu_path():
    """--device cpu runs on the CPU in the reference (infer_two_stage_burr.py:206); here it must fail, not fall through
    to whatever CUDA device is current."""
    from unet_b200 import _lib
    with pytest.raises(_lib.UbError):
        K.resolve_device("cpu")


def test_prefetch_source_keeps_order_and_contract():
    """The decode ring hands out the source's frames in order, then None forever, and releases cleanly mid-stream."""
    frames = np.arange(37 * 4 * 6 * 3, dtype=np.uint8).reshape(37, 4, 6, 3)
    src = K.PrefetchSource(K.ArraySource(frames), depth=5)
    assert (src.total_frames, src.height, src.width) == (37, 4, 6)
    got = []
    while (f := src.read()) is not None:
        got.append(f)
    assert src.read() is None and len(got) == 37 and all(np.array_equal(a, b) for a, b in zip(got, frames))
    src.release()
    src = K.PrefetchSource(K.ArraySource(frames), depth=2)
    assert np.array_equal(src.read(), frames[0])
    src.release()                                   # reader thread blocked on a full queue must stop
    assert not src.t.is_alive()
