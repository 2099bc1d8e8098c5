"""VGG-path pieces on the device (uwie_vgg_* through the C ABI) against the oracle and against the
fixtures generated from the unmodified reference (tests/golden/vgg_*.npz).

Tolerances: the order statistics, the stretch, the dehaze step and the 8-bit resize are exact
(IEEE float32 / integer work); pow differs from torch's CPU pow by a few ulp (bound: 1e-6 absolute,
far inside the 1/255 of north_star); the 18 statistics are summed in float64 instead of NumPy's
pairwise float32 (bound: 1e-5 relative, north_star allows 1e-4)."""
from __future__ import annotations

import glob
import os

import numpy as np
import pytest
import torch

from oracle import restatement as o
from underwater_image_enhancement_b200 import vgg
from underwater_image_enhancement_b200.synth import synth_image

pytestmark = pytest.mark.gpu

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "vgg_*.npz")))
POW_ATOL = 1e-6
FEAT_RTOL = 1e-5


def _nchw(frames):
    return torch.from_numpy(frames.astype(np.float32) / 255.0).permute(0, 3, 1, 2).contiguous().cuda()


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_vgg_path_matches_reference_fixture(path):
    g = np.load(path)
    frame = g["frame"]
    lo, hi, om, ga = (float(x) for x in g["params"])
    img = _nchw(frame[None])
    enh = vgg.DifferentiableEnhancement()
    p = lambda x: torch.tensor([[x]], dtype=torch.float32, device="cuda")       # noqa: E731  (UTM:93-99)
    stretched = enh(img, {"L_low": p(lo), "L_high": p(hi)})[0].cpu().numpy()
    assert np.array_equal(stretched, g["stretched"])
    full = enh(img, {"L_low": p(lo), "L_high": p(hi), "omega": p(om), "gamma": p(ga)})[0].cpu().numpy()
    assert np.abs(full - g["enhanced"]).max() <= POW_ATOL
    feats = vgg.extract_all_features(frame)[0].cpu().numpy()
    assert feats.shape == (79,) and not feats[18:].any()
    assert np.allclose(feats, g["features"], rtol=FEAT_RTOL, atol=1e-7)
    for c in range(3):                                                          # min / max / median are exact
        assert np.array_equal(feats[c * 5 + 2:c * 5 + 5], g["features"][c * 5 + 2:c * 5 + 5])
    pre = vgg.preprocess_for_vgg(frame)[0].cpu().numpy()
    assert np.array_equal(pre, g["preprocessed"])
    pre_f = vgg.preprocess_for_vgg(frame.astype(np.float32) / 255.0)[0].cpu().numpy()   # float input, UTM:41
    assert np.array_equal(pre_f, g["preprocessed"])


@pytest.mark.parametrize("shape", [(33, 250), (97, 131), (720, 1280)])
def test_vgg_enhance_batch_matches_oracle(shape):
    h, w = shape
    frames = np.stack([synth_image(s, h, w) for s in (11, 12, 13)])
    img = _nchw(frames)
    lo, hi = [2.0, 15.0, 29.5], [65.0, 99.0, 80.25]
    om, ga = [0.1, 0.9, 0.55], [0.5, 3.0, 1.3]
    enh = vgg.DifferentiableEnhancement()
    got = enh(img, {"L_low": torch.tensor(lo).view(-1, 1), "L_high": torch.tensor(hi).view(-1, 1),
                    "omega": torch.tensor(om).view(-1, 1), "gamma": torch.tensor(ga).view(-1, 1)}).cpu().numpy()
    ref = o.vgg_enhance(img.cpu().numpy(), lo, hi, om, ga)
    assert np.abs(got - ref).max() <= POW_ATOL
    got_nog = enh(img, {"L_low": torch.tensor(lo), "L_high": torch.tensor(hi), "omega": torch.tensor(om)}).cpu().numpy()
    assert np.array_equal(got_nog, o.vgg_enhance(img.cpu().numpy(), lo, hi, om, None))   # no pow: exact


def test_vgg_enhance_general_floats():
    """Order statistics of arbitrary float32 data (negative values, ties, values above 1)."""
    rng = np.random.default_rng(5)
    x = rng.normal(0.4, 0.5, size=(2, 3, 61, 67)).astype(np.float32)
    x[0, 1, :20] = 0.25                      # heavy ties
    x[1, 2] = np.round(x[1, 2] * 8) / 8      # few distinct values
    lo, hi = [1.0, 30.0], [99.0, 65.0]
    got = vgg.DifferentiableEnhancement()(torch.from_numpy(x).cuda(), {"L_low": lo, "L_high": hi}).cpu().numpy()
    assert np.array_equal(got, o.vgg_enhance(x, lo, hi))


def test_vgg_features_and_preprocess_batch():
    frames = np.stack([synth_image(s, 135, 241) for s in (1, 2, 3, 4)])
    feats = vgg.extract_all_features(frames).cpu().numpy()
    pre = vgg.preprocess_for_vgg(frames, input_size=96).cpu().numpy()
    for i in range(len(frames)):
        assert np.allclose(feats[i], o.vgg_features(frames[i]), rtol=FEAT_RTOL, atol=1e-7)
        assert np.array_equal(pre[i], o.vgg_preprocess(frames[i], 96))
    up = vgg.preprocess_for_vgg(frames[:1], input_size=300).cpu().numpy()        # upscale branch of the resize
    assert np.array_equal(up[0], o.vgg_preprocess(frames[0], 300))


def test_vgg_bad_arguments():
    enh = vgg.DifferentiableEnhancement()
    img = torch.zeros((1, 3, 8, 8), device="cuda")
    with pytest.raises(KeyError):
        enh(img, {"L_low": [1.0]})
    with pytest.raises(ValueError):
        enh(img.cpu(), {"L_low": [1.0], "L_high": [99.0]})
    with pytest.raises(ValueError):
        enh(img, {"L_low": [1.0, 2.0], "L_high": [99.0]})


def test_enhance_from_bytes_equals_enhance_of_float_image(cuda_device):
    """uwie_vgg_enhance_u8 (order statistics from the byte histograms, frame read as bytes) against uwie_vgg_enhance_dev on
    frame / 255 and against the oracle: bit-identical / pow tolerance; the device-side clamps are use_trained_model.py:74-79."""
    import torch
    from underwater_image_enhancement_b200 import vgg
    frames = np.stack([synth_image(90 + i, 135, 241) for i in range(3)])
    d = torch.from_numpy(frames).cuda()
    params = torch.tensor([[0.5, 12.0, 40.0], [99.5, 90.0, 60.0], [0.05, 0.6, 0.95], [0.4, 1.2, 3.5]], device="cuda")
    out8, used8 = vgg.enhance_frames_u8(d, params, clamp_ranges=True)
    # frame.astype(np.float32) / 255.0 on the host, as the reference does (use_trained_model.py:119): torch's CUDA division by
    # a scalar multiplies by the reciprocal and is 1 ulp off for some bytes
    img = torch.from_numpy(np.ascontiguousarray((frames.astype(np.float32) / 255.0).transpose(0, 3, 1, 2))).cuda()
    outf, usedf = vgg.DifferentiableEnhancement().forward_dev(img, params, clamp_ranges=True)
    assert torch.equal(used8, usedf)
    assert torch.equal(used8.cpu(), torch.tensor([[1.0, 12.0, 30.0], [99.0, 90.0, 65.0], [0.1, 0.6, 0.9], [0.5, 1.2, 3.0]]))
    assert torch.equal(out8, outf)
    u = used8.cpu().numpy()
    want = o.vgg_enhance(img.cpu().numpy(), u[0], u[1], u[2], u[3])
    assert np.abs(out8.cpu().numpy() - want).max() <= 1e-6


def test_order_statistics_from_byte_bins_and_the_fallback(cuda_device):
    """uwie_vgg_enhance[_dev] resolves the order statistics in its first pass when a plane has at most one float32 per byte
    bin round(255 x) (k_vgg.cu: vgg_sel_hist_kernel<0>), else by the three-pass radix selection.  Both must give the oracle's
    bits (no gamma: exact) for: bytes / 255 divided on the host (exact quotients), divided by torch on the device (its
    reciprocal multiply is 1 ulp off for some bytes), and a plane where two neighbouring floats share a bin."""
    import torch
    frames = np.stack([synth_image(40 + i, 97, 131) for i in range(2)])
    host = np.ascontiguousarray((frames.astype(np.float32) / 255.0).transpose(0, 3, 1, 2))
    dev = (torch.from_numpy(frames).cuda().to(torch.float32) / 255.0).permute(0, 3, 1, 2).contiguous()
    mixed = host.copy()
    v = mixed[0, 1, 10, 10]
    mixed[0, 1, 10, 11] = np.nextafter(v, np.float32(2.0))          # same byte bin, another float: plane 1 takes the radix passes
    lo, hi, om = [3.0, 12.0], [97.0, 88.0], [0.4, 0.7]
    enh = vgg.DifferentiableEnhancement()
    for name, x in (("host", torch.from_numpy(host).cuda()), ("device", dev), ("mixed", torch.from_numpy(mixed).cuda())):
        got = enh(x, {"L_low": torch.tensor(lo), "L_high": torch.tensor(hi), "omega": torch.tensor(om)}).cpu().numpy()
        want = o.vgg_enhance(x.cpu().numpy(), lo, hi, om, None)
        assert np.array_equal(got, want), name
