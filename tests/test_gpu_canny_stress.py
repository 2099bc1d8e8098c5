"""Exact Canny edge counts on frames built to stress the tiled hysteresis of csrc/k_canny.cu: weak chains that cross many
64x64 tiles before they meet their only strong pixel, spirals, dense textures with many pending components per tile,
frames whose last tiles are partial.  The count is read through feature 65 (Canny density, feature_extraction.py:186-187)
and must equal cv2.Canny(gray, 50, 150) exactly."""
from __future__ import annotations

import cv2
import numpy as np
import pytest

from underwater_image_enhancement_b200 import features as F

pytestmark = pytest.mark.gpu


def _count(gray):
    frame = np.repeat(gray[:, :, None], 3, axis=2)                  # R = G = B: the gray plane is the plane itself
    assert np.array_equal(cv2.cvtColor(frame, cv2.COLOR_RGB2GRAY), gray)
    got = F.extract_all_features_batch(frame[None]).cpu().numpy()[0][65] * gray.size
    want = int(np.count_nonzero(cv2.Canny(gray, 50, 150)))
    return int(round(got)), want


def _wavy_weak_edge(h, w, strong_at):
    """A boundary of step 20 (Sobel magnitude 80: weak) that meanders over the whole frame; a 3-pixel stretch of it has
    step 60 (strong).  Every edge pixel is an edge only through the chain to that stretch."""
    y = np.arange(h)[:, None]
    x = np.arange(w)[None, :]
    bound = w / 2 + (w / 3) * np.sin(y / 23.0)
    img = np.where(x < bound, 100, 120).astype(np.uint8)
    ys = slice(strong_at, strong_at + 3)
    img[ys] = np.where(x < bound[ys], 100, 160).astype(np.uint8)
    return img


def _spiral(n):
    c = n // 2
    t = np.linspace(0, 14 * np.pi, 40000)
    r = 2 + (c - 8) * t / t[-1]
    xs, ys = np.round(c + r * np.cos(t)).astype(int), np.round(c + r * np.sin(t)).astype(int)
    band = np.zeros((n, n), np.uint8)
    band[ys, xs] = 1
    band = cv2.dilate(band, np.ones((3, 3), np.uint8))              # a band three pixels wide ...
    core = np.zeros((n, n), np.uint8)
    core[ys[:60], xs[:60]] = 1
    core = cv2.dilate(core, np.ones((3, 3), np.uint8))
    img = np.full((n, n), 90, np.uint8)
    img[band > 0] = 110                                              # ... whose flanks are weak edges (Sobel magnitude 80)
    img[core > 0] = 170                                              # and strong only at the inner end
    return img


@pytest.mark.parametrize("shape,strong_at", [((256, 320), 5), ((500, 700), 480), ((1080, 1920), 540), ((130, 66), 64)])
def test_weak_chain_across_tiles(cuda_device, shape, strong_at):
    got, want = _count(_wavy_weak_edge(*shape, strong_at))
    assert want > 50 and got == want


def test_weak_chain_without_a_strong_pixel_counts_nothing(cuda_device):
    img = _wavy_weak_edge(300, 400, 10)
    img[10:13] = img[20:23]                                         # remove the strong stretch
    got, want = _count(img)
    assert want == 0 and got == 0


@pytest.mark.parametrize("n", [129, 257, 385])
def test_spiral(cuda_device, n):
    got, want = _count(_spiral(n))
    assert want > 500 and got == want


@pytest.mark.parametrize("seed,amp", [(0, 18), (1, 30), (2, 60)])
def test_dense_texture_many_components_per_tile(cuda_device, seed, amp):
    rng = np.random.default_rng(seed)
    base = rng.integers(0, amp, size=(96, 128)).astype(np.uint8)
    img = cv2.resize(base, (512, 384), interpolation=cv2.INTER_NEAREST) + 60
    img = cv2.GaussianBlur(img, (3, 3), 0)
    got, want = _count(img)
    assert got == want


def test_checkerboards(cuda_device):
    y, x = np.mgrid[0:333, 0:517]
    for period, lo, hi in ((7, 100, 118), (5, 80, 140), (16, 100, 112), (64, 90, 130)):
        img = np.where(((y // period) + (x // period)) % 2 == 0, lo, hi).astype(np.uint8)
        got, want = _count(img)
        assert got == want, (period, got, want)
