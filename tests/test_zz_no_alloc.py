"""uwie_prim_equalize_hist_static: the allocation-free variant of uwie_prim_equalize_hist (scratch in a static
device array of the library, include/uwie.h), called through the C ABI and compared byte for byte with the existing
entry point on consecutive calls.  Collected last on purpose (file name)."""
from __future__ import annotations

import ctypes as C

import numpy as np
import pytest
import torch

from underwater_image_enhancement_b200 import _lib
from underwater_image_enhancement_b200.synth import synth_image

pytestmark = pytest.mark.gpu

MAX_BATCH = 64                                   # UWIE_EQ_STATIC_MAX_BATCH


def _vp(t):
    return C.c_void_p(t.data_ptr())


def _stream(dev):
    return C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


def _existing(lib, frames):
    B, H, W, _ = frames.shape
    out = torch.empty_like(frames)
    ws = torch.empty(B * 3840, dtype=torch.uint8, device=frames.device)
    _lib.check(lib.uwie_prim_equalize_hist(_vp(frames), B, H, W, _vp(out), _vp(ws), ws.numel(), _stream(frames.device)))
    return out


def _static(lib, frames, out):
    B, H, W, _ = frames.shape
    return lib.uwie_prim_equalize_hist_static(_vp(frames), B, H, W, _vp(out), _stream(frames.device))


def test_static_scratch_matches_workspace_entry_point(cuda_device):
    lib = _lib.load()
    dev = torch.device("cuda:0")
    with torch.cuda.device(dev):
        # three calls in a row on one stream, each with other frames and another batch size: a call that saw the
        # previous call's histograms (the static array keeps them) would give other bytes
        cases = [(5, 135, 241, 11), (2, 270, 480, 23), (7, 64, 97, 5)]
        inputs = [torch.from_numpy(np.stack([synth_image(seed + i, h, w) for i in range(b)])).to(dev) for b, h, w, seed in cases]
        wants = [_existing(lib, f) for f in inputs]
        outs = [torch.zeros_like(f) for f in inputs]
        torch.cuda.synchronize()
        for f, out in zip(inputs, outs):
            _lib.check(_static(lib, f, out))
        torch.cuda.synchronize()
        for (b, h, w, _), out, want in zip(cases, outs, wants):
            assert torch.equal(out, want), f"{b}x{h}x{w}: {int((out != want).sum())} bytes differ from uwie_prim_equalize_hist"

        # the same frames again, twice, with no allocation in between: results repeat and free device memory
        # does not shrink (cudaMalloc inside the call would show here)
        free_before, _ = torch.cuda.mem_get_info(dev)
        for _ in range(2):
            outs[0].zero_()
            _lib.check(_static(lib, inputs[0], outs[0]))
            torch.cuda.synchronize()
            assert torch.equal(outs[0], wants[0])
        free_after, _ = torch.cuda.mem_get_info(dev)
        assert free_after >= free_before, f"free device memory fell by {free_before - free_after} bytes across the calls"


def test_static_scratch_matches_cv2(cuda_device):
    import cv2
    lib = _lib.load()
    frames = np.stack([synth_image(31 + i, 120, 200) for i in range(3)])
    t = torch.from_numpy(frames).cuda()
    out = torch.empty_like(t)
    for _ in range(2):
        _lib.check(_static(lib, t, out))
    got = out.cpu().numpy()
    for i in range(len(frames)):
        want = np.stack([cv2.equalizeHist(np.ascontiguousarray(frames[i, :, :, c])) for c in range(3)], axis=2)
        assert np.array_equal(got[i], want)


def test_static_scratch_full_batch_and_limit(cuda_device):
    lib = _lib.load()
    one = synth_image(3, 48, 64)
    frames = torch.from_numpy(np.stack([np.roll(one, i, axis=1) for i in range(MAX_BATCH + 1)])).cuda()
    out = torch.zeros_like(frames)
    # the largest batch the array holds
    _lib.check(_static(lib, frames[:MAX_BATCH], out[:MAX_BATCH]))
    assert torch.equal(out[:MAX_BATCH], _existing(lib, frames[:MAX_BATCH].contiguous()))
    # one more is refused, with a message that names the limit, and writes nothing
    out.zero_()
    rc = _static(lib, frames, out)
    assert rc == -1                                                     # UWIE_ERR_ARG
    msg = lib.uwie_last_error().decode()
    assert "maximum of 64" in msg and "65" in msg, msg
    torch.cuda.synchronize()
    assert int(out.count_nonzero()) == 0
    with pytest.raises(_lib.UwieError, match="maximum of 64"):
        _lib.check(rc)
    # null pointers and empty shapes
    assert lib.uwie_prim_equalize_hist_static(None, 1, 4, 4, _vp(out), None) == -1
    assert _static(lib, frames[:0], out[:0]) == -1
