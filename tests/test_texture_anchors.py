"""Independent anchors for the 22 texture features (FE:80-120), whose arithmetic lives in scikit-image (not installable
here, so the reference fixture for them is made with the oracle's restatement standing in: tests/golden/make_golden_fe.py).
These tests do not go through that fixture:

  * the worked example of scikit-image's own `graycomatrix` documentation (4 x 4 image, 4 grey levels, 4 angles) — its
    four count matrices are re-derived BY HAND in the comments below from the definition "pairs (image[r, c],
    image[r + round(sin a), c + round(cos a)])";
  * `graycoprops` on two matrices whose properties follow from the published formulas by inspection;
  * LBP (P = 8, R = 1, 'uniform') codes derived by hand for a constant image, a vertical step and a plus-shaped pattern,
    with scikit-image's conventions: taps at (-sin, cos) of 2 pi p / 8, bilinear interpolation, zeros outside the image,
    bit = tap >= centre, 'uniform' = at most two changes between the P - 1 ADJACENT pairs, code = number of set bits,
    otherwise P + 1.

The oracle must reproduce them (CPU, below); the CUDA path must reproduce them through uwie_features79
(tests/test_gpu_features.py::test_texture_anchors_on_device).
"""
from __future__ import annotations

import numpy as np

from oracle import restatement as o

DOC_IMAGE = np.array([[0, 0, 1, 1],
                      [0, 0, 1, 1],
                      [0, 2, 2, 2],
                      [2, 2, 3, 3]], dtype=np.uint8)
# angle 0, offset (0, +1): rows give (0,0) (0,1) (1,1) | (0,0) (0,1) (1,1) | (0,2) (2,2) (2,2) | (2,2) (2,3) (3,3)
DOC_0 = np.array([[2, 2, 1, 0], [0, 2, 0, 0], [0, 0, 3, 1], [0, 0, 0, 1]])
# angle pi/4, offset (round(sin a), round(cos a)) = (+1, +1), rows counted downwards:
#   row0->row1 (0,0) (0,1) (1,1) | row1->row2 (0,2) (0,2) (1,2) | row2->row3 (0,2) (2,3) (2,3)
DOC_45 = np.array([[1, 1, 3, 0], [0, 1, 1, 0], [0, 0, 0, 2], [0, 0, 0, 0]])
# angle pi/2, offset (+1, 0): (0,0) (0,0) (1,1) (1,1) | (0,0) (0,2) (1,2) (1,2) | (0,2) (2,2) (2,3) (2,3)
DOC_90 = np.array([[3, 0, 2, 0], [0, 2, 2, 0], [0, 0, 1, 2], [0, 0, 0, 0]])
# angle 3 pi/4, offset (+1, -1): (0,0) (1,0) (1,1) | (0,0) (1,2) (1,2) | (2,2) (2,2) (2,3)
DOC_135 = np.array([[2, 0, 0, 0], [1, 1, 2, 0], [0, 0, 2, 1], [0, 0, 0, 0]])


def glcm_by_definition(img, levels, symmetric=True, normed=True):
    """Straight from the definition, written independently of oracle/restatement.py (loops, no bincount)."""
    offsets = [(0, 1), (1, 1), (1, 0), (1, -1)]             # (round(sin a), round(cos a)) of a = 0, pi/4, pi/2, 3 pi/4
    rows, cols = img.shape
    out = np.zeros((levels, levels, 1, 4))
    for k, (dr, dc) in enumerate(offsets):
        m = np.zeros((levels, levels))
        for r in range(rows):
            for c in range(cols):
                rr, cc = r + dr, c + dc
                if 0 <= rr < rows and 0 <= cc < cols:
                    m[img[r, c], img[rr, cc]] += 1
        if symmetric:
            m = m + m.T
        out[:, :, 0, k] = m / m.sum() if normed else m
    return out


def props_by_definition(P):
    """contrast, dissimilarity, homogeneity, energy, correlation, ASM per angle, from the published formulas."""
    L = P.shape[0]
    res = {k: [] for k in o.GLCM_PROPS}
    for a in range(P.shape[3]):
        p = P[:, :, 0, a]
        con = dis = hom = asm = mi = mj = 0.0
        for i in range(L):
            for j in range(L):
                con += p[i, j] * (i - j) ** 2; dis += p[i, j] * abs(i - j); hom += p[i, j] / (1 + (i - j) ** 2)
                asm += p[i, j] ** 2; mi += i * p[i, j]; mj += j * p[i, j]
        vi = sum(p[i, j] * (i - mi) ** 2 for i in range(L) for j in range(L))
        vj = sum(p[i, j] * (j - mj) ** 2 for i in range(L) for j in range(L))
        cov = sum(p[i, j] * (i - mi) * (j - mj) for i in range(L) for j in range(L))
        cor = 1.0 if (vi ** 0.5 < 1e-15 or vj ** 0.5 < 1e-15) else cov / (vi ** 0.5 * vj ** 0.5)
        for k, v in zip(o.GLCM_PROPS, (con, dis, hom, asm ** 0.5, cor, asm)):
            res[k].append(v)
    return res


def test_definition_helper_reproduces_the_documented_matrices():
    raw = glcm_by_definition(DOC_IMAGE, 4, symmetric=False, normed=False)
    for k, want in enumerate((DOC_0, DOC_45, DOC_90, DOC_135)):
        assert np.array_equal(raw[:, :, 0, k], want), k


def test_oracle_graycomatrix_on_the_documented_example():
    """The oracle restates the symmetric, normalised form (FE:107-112): (M + M^T) / sum for each documented matrix.
    A symmetric matrix does not care whether the row axis of the offset points up or down, which is why both
    conventions give the same features."""
    got = o.sk_graycomatrix(DOC_IMAGE, levels=4)
    assert got.shape == (4, 4, 1, 4)
    for k, m in enumerate((DOC_0, DOC_45, DOC_90, DOC_135)):
        want = (m + m.T) / (m + m.T).sum()
        assert np.allclose(got[:, :, 0, k], want, atol=0, rtol=1e-15), k


def test_oracle_graycoprops_on_matrices_with_known_properties():
    P = np.zeros((2, 2, 1, 2))
    P[:, :, 0, 0] = [[0.5, 0.0], [0.0, 0.5]]      # only equal pairs
    P[:, :, 0, 1] = [[0.0, 0.5], [0.5, 0.0]]      # only (0, 1) / (1, 0) pairs
    want = {"contrast": [0, 1], "dissimilarity": [0, 1], "homogeneity": [1, 0.5], "ASM": [0.5, 0.5],
            "energy": [0.5 ** 0.5, 0.5 ** 0.5], "correlation": [1, -1]}
    for prop, v in want.items():
        assert np.allclose(o.sk_graycoprops(P, prop).reshape(-1), v, atol=1e-15), prop
    rng = np.random.default_rng(0)
    img = rng.integers(0, 6, (9, 11))
    Pd = glcm_by_definition(img, 6)
    assert np.allclose(o.sk_graycomatrix(img, 6), Pd, rtol=1e-14)
    for prop, v in props_by_definition(Pd).items():
        assert np.allclose(o.sk_graycoprops(Pd, prop).reshape(-1), v, rtol=1e-12, atol=1e-15), prop


def lbp_expected_constant(h, w, positive=True):
    """Constant image.  value 0: every tap is 0 >= 0 -> all 8 bits -> code 8 everywhere.  value > 0: interior 8;
    on an edge the three taps towards the outside are 0 or a bilinear mix with 0 (< centre): bits 1,0,0,0,1,1,1,1 up to
    rotation -> 5 ones, 2 adjacent changes -> code 5; in a corner 1,0,0,0,0,0,1,1 up to rotation -> code 3."""
    out = np.full((h, w), 8.0)
    if positive:
        out[0, :] = out[-1, :] = out[:, 0] = out[:, -1] = 5.0
        out[0, 0] = out[0, -1] = out[-1, 0] = out[-1, -1] = 3.0
    return out


def test_oracle_lbp_hand_derived_codes():
    assert np.array_equal(o.sk_local_binary_pattern_uniform(np.zeros((6, 7), np.uint8)), lbp_expected_constant(6, 7, False))
    assert np.array_equal(o.sk_local_binary_pattern_uniform(np.full((6, 7), 90, np.uint8)), lbp_expected_constant(6, 7, True))
    # vertical step 0 | 10 at column 4: on the bright side next to the step the W, NW, SW taps are darker:
    # E 1, NE 1, N 1, NW 0 (10 * 0.293 + 0 * 0.707 < 10), W 0, SW 0, S 1, SE 1 -> 5 ones, 2 changes -> code 5
    step = np.zeros((7, 8), np.uint8)
    step[:, 4:] = 10
    lbp = o.sk_local_binary_pattern_uniform(step)
    assert np.all(lbp[1:-1, 4] == 5) and np.all(lbp[1:-1, 5:-1] == 8) and np.all(lbp[:, :4] == 8)
    # plus pattern: centre 5, N S E W = 10, diagonals 0: diagonal taps = .0858 * 5 + 2 * .2071 * 10 = 4.571 < 5
    # -> bits 1,0,1,0,1,0,1,0 -> 7 changes -> not uniform -> P + 1 = 9
    plus = np.zeros((5, 5), np.uint8)
    plus[2, 2] = 5
    plus[1, 2] = plus[3, 2] = plus[2, 1] = plus[2, 3] = 10
    assert o.sk_local_binary_pattern_uniform(plus)[2, 2] == 9
    # a single bright pixel: every tap is darker -> code 0
    dot = np.zeros((5, 5), np.uint8)
    dot[2, 2] = 200
    assert o.sk_local_binary_pattern_uniform(dot)[2, 2] == 0
