"""ub_sobel_edges_static: the allocation-free variant of ub_sobel_edges (static device scratch, no context).

Collected last on purpose.  Called through the C ABI exactly as a foreign caller would:
plain device pointers and sizes.  The result must equal the existing entry point's bit for bit (integer path), on two
calls in a row -- the second on frames with SMALLER gradient maxima, so stale scratch contents would change the answer."""
import ctypes as C

import numpy as np
import pytest
import torch

import unet_b200.synth as S

pytestmark = pytest.mark.gpu


def _static_call(lib, gray, thresh):
    out = torch.empty_like(gray)
    b, h, w = gray.shape
    rc = lib.ub_sobel_edges_static(C.c_void_p(gray.data_ptr()), b, h, w, thresh, C.c_void_p(out.data_ptr()),
                                   C.c_void_p(torch.cuda.current_stream().cuda_stream))
    return rc, out


def test_sobel_edges_static_matches_context_entry_point():
    from unet_b200 import _lib
    from oracle import cvnp                                   # checker only
    lib = _lib.lib()
    torch.cuda.set_device(0)
    ctx = _lib.Context(0, max_batch=3, frame_hw=(448, 800), net_hw=(16, 16))
    frames = np.concatenate([S.frames_cable(2, 448, 800, seed=21), S.frames_speckle(1, 448, 800, seed=22)])
    gray1 = torch.from_numpy(np.ascontiguousarray(frames[..., 1])).cuda()
    gray2 = (gray1 // 3).contiguous()                         # every per-frame gradient maximum shrinks
    want1, want2 = ctx.sobel_edges(gray1), ctx.sobel_edges(gray2)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    rc1, got1 = _static_call(lib, gray1, 50)
    rc2, got2 = _static_call(lib, gray2, 50)
    torch.cuda.synchronize()
    assert rc1 == 0 and rc2 == 0, lib.ub_last_error(None)
    assert torch.equal(got1, want1) and torch.equal(got2, want2)
    assert int(got1.sum()) > 0 and not torch.equal(got1, got2)
    # and against the oracle's restatement of the reference statements (infer_enhanced_burr.py:93-97)
    assert np.array_equal(got2[2].cpu().numpy(), cvnp.sobel_edges(gray2[2].cpu().numpy()))
    # a ragged shape, again twice
    g3 = gray1[:1, :37, :53].contiguous()
    for _ in range(2):
        rc, got = _static_call(lib, g3, 50)
        assert rc == 0 and torch.equal(got, ctx.sobel_edges(g3))
    ctx.close()


def test_sobel_edges_static_rejects_oversized_batch_and_null():
    from unet_b200 import _lib
    lib = _lib.lib()
    torch.cuda.set_device(0)
    dummy = torch.zeros(16, dtype=torch.uint8, device="cuda")
    p = C.c_void_p(dummy.data_ptr())
    rc = lib.ub_sobel_edges_static(p, 4097, 1, 1, 50, p, None)            # UB_STATIC_MAX_FRAMES = 4096
    assert rc == -1
    assert b"UB_STATIC_MAX_FRAMES" in lib.ub_last_error(None)
    assert lib.ub_sobel_edges_static(None, 1, 4, 4, 50, p, None) == -1
    rc, out = _static_call(lib, dummy.view(1, 4, 4), 50)                   # still usable afterwards
    torch.cuda.synchronize()
    assert rc == 0 and int(out.sum()) == 0
