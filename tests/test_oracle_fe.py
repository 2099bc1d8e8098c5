"""The FE restatement (oracle/restatement.py: fe_*) against the fixture generated from the unmodified
reference FeatureExtractor (tests/golden/make_golden_fe.py).  CPU only."""
from __future__ import annotations

import os

import numpy as np

from oracle import restatement as o
from underwater_image_enhancement_b200.synth import synth_image

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "fe_features.npz"))


def test_fe_oracle_reproduces_reference_bit_for_bit():
    for (seed, h, w), want in zip(G["cases"], G["features"]):
        got = o.fe_all(synth_image(int(seed), int(h), int(w)))
        assert got.shape == (79,)
        assert np.array_equal(got, want), (seed, np.nonzero(got != want))


def test_group_sizes():
    assert [n for _, _, n in o.FE_GROUPS] == [35, 22, 5, 7, 10]        # FE:249-297 -> 79
    img = synth_image(2, 64, 80).astype(np.float32) / 255.0
    for _, fn, n in o.FE_GROUPS:
        assert len(fn(img)) == n


def test_lbp_uniform_codes_on_known_patterns():
    flat = np.full((5, 5), 7, np.uint8)
    codes = o.sk_local_binary_pattern_uniform(flat)
    assert codes[2, 2] == 8                    # all eight taps >= centre
    spike = np.zeros((5, 5), np.uint8); spike[2, 2] = 9
    assert o.sk_local_binary_pattern_uniform(spike)[2, 2] == 0


def test_glcm_is_symmetric_and_normalised():
    g = (np.random.default_rng(1).random((16, 16)) * 255).astype(np.uint8)
    P = o.sk_graycomatrix(g)
    assert np.allclose(P.sum(axis=(0, 1)), 1.0)
    assert np.array_equal(P[:, :, 0, 1], P[:, :, 0, 1].T)
