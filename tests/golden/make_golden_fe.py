"""Generate tests/golden/fe_features.npz by running the UNMODIFIED reference FeatureExtractor here.

    python tests/golden/make_golden_fe.py         # needs /root/reference (read-only mount)

feature_extraction.py imports scikit-image (not installable here).  The stand-in module gives it
shannon_entropy (three lines, see make_golden.py) and, for the texture group only, the oracle's
restatements of local_binary_pattern / graycomatrix / graycoprops — so the colour, frequency, edge and
quality groups (57 values) are pinned to the reference's own code path, and the texture group
(22 values) is pinned to the reference's CALL SITES (FE:93-118) but not to scikit-image itself
("parity unpinned" for those 22).
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

from make_golden import REF, install_skimage_standin  # noqa: E402

CASES = ((0, 120, 160), (1, 120, 160), (3, 135, 241), (7, 250, 333), (5, 256, 256), (8, 360, 640))


def main():
    from oracle import restatement as o
    from underwater_image_enhancement_b200.synth import synth_image
    install_skimage_standin()
    import skimage
    # The stand-ins restate ONE configuration of the scikit-image functions — the one feature_extraction.py:97, 107-112
    # asks for — and refuse anything else instead of silently ignoring the arguments.
    def local_binary_pattern(image, P, R, method="default"):
        assert (P, R, method) == (8, 1, "uniform"), (P, R, method)
        return o.sk_local_binary_pattern_uniform(image, P, R)

    def graycomatrix(image, distances, angles, levels=None, symmetric=False, normed=False):
        assert list(distances) == [1] and levels == 256 and symmetric is True and normed is True, (distances, levels, symmetric, normed)
        assert np.allclose(list(angles), [0, np.pi / 4, np.pi / 2, 3 * np.pi / 4]), angles
        return o.sk_graycomatrix(image, levels)

    skimage.feature.local_binary_pattern = local_binary_pattern
    skimage.feature.graycomatrix = graycomatrix
    skimage.feature.graycoprops = o.sk_graycoprops
    sys.path.insert(0, REF)
    import feature_extraction as fe
    feats = []
    for seed, h, w in CASES:
        img = synth_image(seed, h, w).astype(np.float32) / 255.0          # main.py:108
        f = np.asarray(fe.FeatureExtractor.extract_all_features(img), dtype=np.float64)
        assert f.shape == (79,), f.shape
        feats.append(f)
    path = os.path.join(HERE, "fe_features.npz")
    np.savez_compressed(path, cases=np.asarray(CASES, np.int64), features=np.stack(feats))
    print(path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
