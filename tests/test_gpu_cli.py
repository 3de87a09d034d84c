"""GPU: the drop-in command lines end to end (SURVEY 8b "CLI contract").  The printed progress
lines must be the reference's lines for the statistics the device produced; the statistics are
checked against the oracle's per-frame step (argmax tolerance for the masks, bit-exact stage 2
given the device's own cable mask is covered in test_gpu_net / test_gpu_stage2)."""
import re

import numpy as np
import pytest
import torch

import unet_b200.synth as S
from oracle import ref_pipeline as R
from unet_b200 import cli_common as K
from unet_b200 import cli_enhanced, cli_two_stage

pytestmark = pytest.mark.gpu
ARGMAX_MIN_AGREE = 0.999


@pytest.fixture(scope="module")
def sd():
    return S.make_state_dict(seed=0)


def test_two_stage_cli_synthetic(tmp_path, sd):
    ckpt = tmp_path / "best_model.pth"
    torch.save({"model": sd}, ckpt)                                  # exercises the reference's loader path (:215-216)
    lines = []
    rec = cli_two_stage.main(["--synthetic", "70", "--model", str(ckpt), "--output", str(tmp_path / "out"),
                              "--batch", "32", "--print-interval", "30", "--no-video-out"], out=lines.append)
    assert rec.shape == (70, 5)
    text = "\n".join(lines)
    assert "两阶段毛刺检测系统" in text and "固定ROI（800x448）: X[218, 421] Y[0, 448]" in text
    assert "处理分辨率: 800x448 @ 30.00fps, 总帧数: 70" in text
    prog = [l for l in lines if re.match(r"^\[ *\d+\] Frame", l)]
    roi = (218, 0, 421, 448)
    assert prog == [K.progress_line(int(r[0]), int(r[1]), 70, int(r[2]), int(r[3]), int(r[4]), roi) for r in rec if r[0] % 30 == 0]
    assert len(prog) == 2
    assert any(l.startswith("检测到毛刺的帧数:") for l in lines)
    frames = S.frames_cable(70, 448, 800, seed=0)
    for idx in (0, 29, 69):
        ref = R.frame_step(frames[idx], sd, "medium")
        area = 448 * 800
        assert abs(int(rec[idx, 2]) - ref["cable_px"]) <= (1 - ARGMAX_MIN_AGREE) * area
        assert abs(int(rec[idx, 3]) - ref["tape_px"]) <= (1 - ARGMAX_MIN_AGREE) * area


def test_two_stage_cli_video_file_rotate_normalize_and_output(tmp_path, sd):
    """A real (mp4v) file through cv2.VideoCapture, --rotate --normalize-resolution --frame-stride 2,
    and result.mp4 written with the overlay."""
    import cv2
    src = S.frames_cable(9, 270, 480, seed=3)                         # landscape; rotated -> 480x270... then 800x448
    path = str(tmp_path / "in.mp4")
    w = cv2.VideoWriter(path, cv2.VideoWriter_fourcc(*"mp4v"), 25.0, (480, 270))
    if not w.isOpened():
        pytest.skip("this cv2 build cannot encode mp4v")
    for f in src:
        w.write(f)
    w.release()
    lines = []
    rec = cli_two_stage.main(["--video", path, "--output", str(tmp_path / "o"), "--batch", "4", "--frame-stride", "2",
                              "--print-interval", "2", "--rotate", "--normalize-resolution"], out=lines.append, state_dict=sd)
    text = "\n".join(lines)
    assert "原始视频: 480x270" in text and "旋转后: 270x480" in text and "归一化到: 800x448" in text
    assert rec.shape[0] == 4 and rec[:, 1].tolist() == [2, 4, 6, 8]
    cap = cv2.VideoCapture(str(tmp_path / "o" / "result.mp4"))
    assert cap.isOpened() and int(cap.get(cv2.CAP_PROP_FRAME_COUNT)) == 4
    assert (int(cap.get(cv2.CAP_PROP_FRAME_WIDTH)), int(cap.get(cv2.CAP_PROP_FRAME_HEIGHT))) == (800, 448)
    # statistics equal the reference loop's statements on the decoded frames (decoder output is what both see)
    cap_in = cv2.VideoCapture(path)
    k = 0
    for i in range(1, 10):
        ok, fr = cap_in.read()
        assert ok
        if i % 2:
            continue
        fr = R.normalize_frame(fr, rotate=True, size=(800, 448))
        ref = R.frame_step(fr, sd, "medium")
        assert abs(int(rec[k, 2]) - ref["cable_px"]) <= (1 - ARGMAX_MIN_AGREE) * 448 * 800
        k += 1


def test_unopenable_video_prints_error_and_returns(tmp_path, sd):
    lines = []
    assert cli_two_stage.main(["--video", str(tmp_path / "nope.mp4"), "--output", str(tmp_path / "o")],
                              out=lines.append, state_dict=sd) == 0
    assert lines[-1] == f"错误: 无法打开视频 {tmp_path / 'nope.mp4'}"


@pytest.mark.parametrize("enhance", [False, True])
def test_enhanced_cli(tmp_path, sd, enhance):
    lines = []
    argv = ["--synthetic", "6", "--synthetic-size", "400x224", "--output", str(tmp_path / "e"), "--batch", "4", "--no-video-out"]
    rec = cli_enhanced.main(argv + (["--enhance"] if enhance else []), out=lines.append, state_dict=sd)
    text = "\n".join(lines)
    assert f"图像增强: {'启用' if enhance else '禁用'}" in text and "ROI: X[200, 600] Y[0, 448]" in text
    assert ("图像增强: CLAHE + 去噪 + 锐化" in text) == enhance
    assert "原始视频: 224x400 @ 30.00fps, 总帧数: 6" in text
    assert rec.shape == (6, 5)
    frames = S.frames_cable(6, 400, 224, seed=0)
    for idx in (0, 5):
        fr = R.normalize_frame(frames[idx], rotate=True, size=(800, 448))
        ref = R.frame_step(fr, sd, variant="enhanced", do_enhance=enhance)
        assert abs(int(rec[idx, 2]) - ref["cable_px"]) <= (1 - ARGMAX_MIN_AGREE) * 448 * 800
        assert abs(int(rec[idx, 3]) - ref["tape_px"]) <= (1 - ARGMAX_MIN_AGREE) * 448 * 800
