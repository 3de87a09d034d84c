"""CPU-only checks: the C-ABI library loads and exports every symbol the header declares,
host-side logic (ROI mapping, sharding, statistics formatting, module key layout), and the
world_size-2 statistics gather over gloo.  No compute call is made without a GPU."""
import os
import re
import socket
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from unet_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        _lib.build()
    hdr = open(os.path.join(ROOT, "include", "unetb200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(ub_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    l = _lib.lib()
    for name in declared:
        assert hasattr(l, name), name
    assert l.ub_version() >= 100
    # argument validation works without a GPU and never crashes
    assert l.ub_forward(None, None, 1, 16, 16, None, None) < 0
    assert b"null context" in l.ub_last_error(None)


def test_no_cpu_fallback():
    from unet_b200 import _lib, NestedUNet
    m = NestedUNet(3).eval()
    if not torch.cuda.is_available():
        with pytest.raises(_lib.UbError):
            m(torch.zeros(1, 3, 16, 16))
        with pytest.raises(_lib.UbError):
            _lib.Context(0)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "unet-_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src, f


def test_module_state_dict_layout(golden_net):
    from unet_b200 import NestedUNet
    import unet_b200.synth as S
    m = NestedUNet(num_classes=3, deep_supervision=True, pretrained_encoder=False)
    assert list(m.state_dict().keys()) == list(golden_net["state_dict_keys"])
    assert sum(p.numel() for p in m.parameters()) == 7853964
    m.load_state_dict(S.make_state_dict(0), strict=True)
    m2 = NestedUNet(num_classes=3, deep_supervision=False)
    assert len(m2.state_dict()) == 128


def test_roi_and_stats_line():
    from unet_b200.ops import map_roi_to_original
    from unet_b200.pipeline import shard_range, stats_line
    from oracle import ref_pipeline as R
    for wh in ((800, 448), (1920, 1080), (640, 480), (512, 512)):
        assert map_roi_to_original(wh) == R.map_roi(wh)
    assert map_roi_to_original((800, 448)) == (218, 0, 421, 448)
    roi = (218, 0, 421, 448)
    for args in ((1, 1, 10, 19319, 71625, 0), (60, 60, 900, 0, 0, 12), (7, 14, 20, 90944, 1, 1)):
        assert stats_line(*args, roi) == R.stats_line(*args, roi)
    n, world = 1003, 8
    cover = []
    for r in range(world):
        a, b = shard_range(n, r, world)
        cover += list(range(a, b))
    assert cover == list(range(n))


def test_synth_is_deterministic():
    import unet_b200.synth as S
    assert np.array_equal(S.frames_cable(1, 64, 96, seed=3), S.frames_cable(1, 64, 96, seed=3))
    a, b = S.make_state_dict(0), S.make_state_dict(0)
    assert all(torch.equal(a[k], b[k]) for k in a)


_WORKER = r"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from unet_b200.pipeline import InspectionPipeline, shard_range
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo", rank=rank, world_size=world)
n = 11
a, b = shard_range(n, rank, world)
local = torch.stack([torch.arange(b - a, dtype=torch.int32), torch.arange(a, b, dtype=torch.int32) * 10,
                     torch.arange(a, b, dtype=torch.int32) * 100, (torch.arange(a, b, dtype=torch.int32) % 3)], 1)
out = InspectionPipeline.gather_stats(local, a)
want = torch.stack([torch.arange(n), torch.arange(n) * 10, torch.arange(n) * 100, torch.arange(n) % 3], 1).int()
assert torch.equal(out, want), (rank, out)
# equal shares: one collective, no size exchange
m = 6
loc = torch.stack([torch.arange(m, dtype=torch.int32), torch.full((m,), rank, dtype=torch.int32),
                   torch.arange(m, dtype=torch.int32) + 7, torch.zeros(m, dtype=torch.int32)], 1)
out = InspectionPipeline.gather_stats(loc, rank * m, equal_counts=True)
assert out.shape == (world * m, 4) and out[:, 0].tolist() == list(range(world * m))
assert out[:, 1].tolist() == [r for r in range(world) for _ in range(m)]
dist.destroy_process_group()
print("ok", rank)
"""


def test_gather_stats_gloo_world2(tmp_path):
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
        out, _ = p.communicate(timeout=120)
        assert p.returncode == 0, out


def test_contour_stamps_match_installed_opencv():
    """The stamps hard-coded in csrc/annotate.cu are what cv2.line paints for a unit step / a single point (thickness 2, 3)."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_stamps", os.path.join(ROOT, "tools", "make_stamps.py"))
    ms = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ms)
    from unet_b200 import _lib
    for thickness in (2, 3):
        for code in range(9):
            assert _lib.lib().ub_contour_stamp(thickness, code) == ms.stamp(thickness, code), (thickness, code)
    assert _lib.lib().ub_contour_stamp(4, 0) == 0 and _lib.lib().ub_contour_stamp(2, 9) == 0
