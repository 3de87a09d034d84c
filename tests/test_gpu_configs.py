"""Config-shaped parity: BASELINE.json configs 4 and 5 run as ONE call each through the pipeline entry points and are
checked against the oracle, and the status the CLI prints per frame ([BURR!] / [OK], infer_two_stage_burr.py:333-343)
is compared between the GPU pipeline and the all-fp32 oracle over 64 synthetic frames."""
import cv2
import numpy as np
import pytest
import torch

import unet_b200.synth as S
from oracle import ref_pipeline as R

pytestmark = pytest.mark.gpu

ARGMAX_MIN_AGREE = 0.999   # BASELINE.json north_star: "argmax mask agreeing on >= 99.9 % of pixels"


@pytest.fixture(scope="module")
def sd():
    return S.make_state_dict(0)


def status_agreement(pipe, frames, sd):
    """(agreeing, total, rows): rows hold (gpu burr_px, oracle burr_px) per frame."""
    stats = np.concatenate([pipe.run(torch.from_numpy(frames[i:i + pipe.batch]).cuda())[0].cpu().numpy()
                            for i in range(0, len(frames), pipe.batch)])
    rows = [(int(stats[i, 3]), R.frame_step(frames[i], sd, "medium")["burr_px"]) for i in range(len(frames))]
    return sum((g > 0) == (o > 0) for g, o in rows), len(rows), rows


def test_config5_pipeline_512x512_vs_oracle(sd):
    """BASELINE config 5: 512x512 frames (no resize in preprocess, ROI x 140..270 over the full height)."""
    from unet_b200.pipeline import InspectionPipeline
    frames = np.concatenate([S.frames_cable(2, 512, 512, seed=14), S.frames_speckle(1, 512, 512, seed=15)])
    pipe = InspectionPipeline(sd, batch=3, frame_hw=(512, 512))
    assert tuple(pipe.roi) == R.map_roi((512, 512)) == (140, 0, 270, 512)
    stats, masks = pipe.run(torch.from_numpy(frames).cuda(), want_masks=True)
    stats, masks = stats.cpu().numpy(), masks.cpu().numpy()
    assert np.array_equal(stats, pipe.run_host(torch.from_numpy(frames).pin_memory()).numpy())
    for i in range(3):
        ref = R.frame_step(frames[i], sd, "medium")
        assert ref["roi"] == tuple(pipe.roi)
        agree_c = (masks[i, 0] == ref["mask_cable"]).mean()
        agree_t = (masks[i, 1] == ref["mask_tape"]).mean()
        print(f"config 5 frame {i}: stats {stats[i].tolist()} oracle {[ref['cable_px'], ref['tape_px'], ref['burr_px']]} "
              f"mask agreement cable {agree_c:.5f} tape {agree_t:.5f}")
        assert agree_c >= ARGMAX_MIN_AGREE and agree_t >= ARGMAX_MIN_AGREE
        gray = cv2.cvtColor(frames[i], cv2.COLOR_BGR2GRAY)
        burr = R.burrs_standard(gray, masks[i, 0], R.BURR_CONFIGS["medium"])
        assert np.array_equal(masks[i, 2], burr)                 # integer stage: bit-exact on the same cable mask
        assert stats[i].tolist() == [i, int(masks[i, 0].sum()), int(masks[i, 1].sum()), int(burr.sum())]
    pipe.ctx.close()


def test_config4_full_chain_one_call_vs_oracle(sd):
    """BASELINE config 4 in one call: 1080x1920 source -> rotate 90 CCW -> 800x448 -> enhance_image -> net -> enhanced
    multi-scale detector (infer_enhanced_burr.py:273-312), through ub_pipeline_submit_src with HOST frames."""
    from unet_b200.pipeline import InspectionPipeline
    src = S.frames_cable(2, 1080, 1920, seed=21)
    pipe = InspectionPipeline(sd, batch=2, frame_hw=(448, 800), variant="enhanced", enhance=True)
    host = torch.from_numpy(src).pin_memory()
    got = [s.numpy() for s in pipe.stream_host(iter([host, host[:1]]), src_rotate=True)]
    assert got[0].shape == (2, 4) and got[1].shape == (1, 4) and np.array_equal(got[1][0], got[0][0])
    # the same chain step by step on the device, with masks
    norm = pipe.ctx.normalize_frames(torch.from_numpy(src).cuda(), (800, 448), rotate=True)
    stats, masks = pipe.run(norm, want_masks=True)
    stats, masks = stats.cpu().numpy(), masks.cpu().numpy()
    assert np.array_equal(stats, got[0])
    for i in range(2):
        fr = R.normalize_frame(src[i], rotate=True, size=(800, 448))
        assert np.array_equal(norm[i].cpu().numpy(), fr)
        ref = R.frame_step(fr, sd, variant="enhanced", do_enhance=True)
        agree_c = (masks[i, 0] == ref["mask_cable"]).mean()
        print(f"config 4 frame {i}: stats {stats[i].tolist()} oracle {[ref['cable_px'], ref['tape_px'], ref['burr_px']]} "
              f"cable mask agreement {agree_c:.5f}")
        assert agree_c >= ARGMAX_MIN_AGREE
        gray = cv2.cvtColor(R.enhance(fr), cv2.COLOR_BGR2GRAY)    # the detector sees the ENHANCED frame (:279-280,:311)
        burr = R.burrs_enhanced(gray, masks[i, 0], R.ENHANCED_CONFIG)
        assert np.array_equal(masks[i, 2], burr)
        assert stats[i].tolist() == [i, int(masks[i, 0].sum()), int(masks[i, 1].sum()), int(burr.sum())]
    pipe.ctx.close()


def test_status_agreement_over_64_frames(sd):
    """How often the 16-bit network flips what the CLI prints: [BURR!] iff burr_px > 0."""
    from unet_b200.pipeline import InspectionPipeline
    torch.set_num_threads(min(16, torch.get_num_threads()))
    frames = S.status_frames()
    pipe = InspectionPipeline(sd, batch=16)
    ok, n, rows = status_agreement(pipe, frames, sd)
    burr = sum(o > 0 for _, o in rows)
    rel = [abs(g - o) / max(o, 1) for g, o in rows]
    print(f"status agreement {ok}/{n} ({burr} [BURR!] / {n - burr} [OK] by the oracle); "
          f"burr_px relative difference: median {np.median(rel):.4f} max {max(rel):.4f}")
    assert n == 64 and ok == n
    pipe.ctx.close()


def test_graph_replay_equals_plain_step(sd):
    """ub_pipeline_graph (first call plain, second captures, later calls replay) gives the statistics of ub_pipeline for
    every refill of the same frame tensor, at the small batches it exists for."""
    from unet_b200.pipeline import InspectionPipeline
    for b in (1, 4):
        pipe = InspectionPipeline(sd, batch=b)
        frames = torch.empty((b, 448, 800, 3), dtype=torch.uint8, device="cuda")
        side = torch.cuda.Stream()
        with torch.cuda.stream(side):                      # a capturable stream (not the legacy default)
            for rep in range(5):
                fr = np.concatenate([S.frames_cable(b, 448, 800, seed=60 + rep)[: max(1, b - 1)],
                                     S.frames_speckle(1, 448, 800, seed=70 + rep)])[:b]
                frames.copy_(torch.from_numpy(fr))
                want = pipe.run(frames)[0].cpu().numpy()
                n0 = pipe.ctx.launch_count()
                got = pipe.run_graph(frames).cpu().numpy()
                assert np.array_equal(got, want), (b, rep)
                assert pipe.ctx.launch_count() - n0 > 20    # the replay accounts for the kernels it launches
        side.synchronize()
        pipe.ctx.close()
