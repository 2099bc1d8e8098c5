"""The VGG-path restatement (oracle/restatement.py: vgg_*) against fixtures generated from the
unmodified reference (tests/golden/make_golden_vgg.py).  CPU only."""
from __future__ import annotations

import glob
import os

import numpy as np
import pytest

from oracle import restatement as o

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "vgg_*.npz")))


def test_fixtures_present():
    assert len(GOLDEN) >= 4


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_vgg_oracle_matches_reference(path):
    g = np.load(path)
    frame = g["frame"]
    lo, hi, om, ga = (float(x) for x in g["params"])
    img = (frame.astype(np.float32) / 255.0).transpose(2, 0, 1)[None]
    full = o.vgg_enhance(img, [lo], [hi], [om], [ga])[0]
    # stretch and dehaze are IEEE float32 operations (bit-identical); torch's and NumPy's float32
    # pow may differ in the last place
    assert np.array_equal(o.vgg_enhance(img, [lo], [hi])[0], g["stretched"])
    assert np.abs(full - g["enhanced"]).max() <= 2.4e-7
    assert np.array_equal(o.vgg_features(frame), g["features"])
    assert np.array_equal(o.vgg_preprocess(frame), g["preprocessed"])


def test_vgg_rank_follows_python_float_arithmetic():
    assert o.vgg_stretch_rank(10.0, 19200) == 1920
    assert o.vgg_stretch_rank(100.0, 19200) == 19199          # clamped (VGG:81-82)
    assert o.vgg_stretch_rank(0.0, 7) == 0
    # a float32 level is widened to double before the division (tensor.item())
    assert o.vgg_stretch_rank(np.float32(97.25), 135 * 241) == int((97.25 / 100.0) * 135 * 241)
