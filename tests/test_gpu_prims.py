"""Per-primitive operators (uwie_prim_*) against cv2 / NumPy / the oracle.

Integer primitives are bit-exact (CLAHE on a plane, equalizeHist, quantise, the float32 transmission head on
bytes).  Float-image primitives (arbitrary float32 / float64 images, not byte-valued) are held to north_star's
float-stage bar and mostly far inside it: their box sums and Q statistics are float64 sums in another order
than cv2's / NumPy's."""
from __future__ import annotations

import cv2
import numpy as np
import pytest
import torch

from oracle import restatement as o
from underwater_image_enhancement_b200 import prims
from underwater_image_enhancement_b200.synth import synth_image

pytestmark = pytest.mark.gpu


def _np(t):
    return t.cpu().numpy()


# ------------------------------------------------------------------ integer primitives: exact
@pytest.mark.parametrize("h,w,grid,clip", [(64, 64, (8, 8), 2.0), (135, 241, (8, 8), 3.0), (250, 333, (4, 6), 1.5),
                                            (1080, 1920, (8, 8), 2.0), (2160, 3840, (8, 8), 4.0), (97, 1300, (2, 2), 40.0)])
def test_clahe_lab_bit_exact(cuda_device, h, w, grid, clip):
    frames = np.stack([synth_image(3 + i, h, w) for i in range(2)])
    planes = np.stack([cv2.cvtColor(f, cv2.COLOR_RGB2LAB)[:, :, 0] for f in frames])
    got = _np(prims.clahe_lab(planes, clip, grid))
    op = cv2.createCLAHE(clipLimit=clip, tileGridSize=grid)
    for i in range(len(planes)):
        want = op.apply(np.ascontiguousarray(planes[i]))
        assert np.array_equal(got[i], want), f"plane {i}: {np.count_nonzero(got[i] != want)} bytes differ"


def test_clahe_lab_flat_and_extreme_planes(cuda_device):
    h, w = 120, 200
    rng = np.random.default_rng(2)
    planes = np.stack([np.full((h, w), 77, np.uint8), rng.integers(0, 256, (h, w), dtype=np.uint8),
                       (np.arange(h * w).reshape(h, w) % 2 * 255).astype(np.uint8)])
    got = _np(prims.clahe_lab(planes, 2.0, (8, 8)))
    op = cv2.createCLAHE(clipLimit=2.0, tileGridSize=(8, 8))
    for i in range(len(planes)):
        assert np.array_equal(got[i], op.apply(np.ascontiguousarray(planes[i]))), i


@pytest.mark.parametrize("h,w", [(96, 128), (135, 241), (1080, 1920)])
def test_equalize_hist_bit_exact(cuda_device, h, w):
    frames = np.stack([synth_image(7, h, w), np.full((h, w, 3), 9, np.uint8)])
    got = _np(prims.equalize_hist(frames))
    for i in range(2):
        for c in range(3):
            assert np.array_equal(got[i, :, :, c], cv2.equalizeHist(np.ascontiguousarray(frames[i, :, :, c])))


@pytest.mark.parametrize("h,w,clip", [(1080, 1920, True), (2160, 3840, False), (33, 251, True), (5, 7, False)])
def test_transmission_head_u8_bit_exact(cuda_device, h, w, clip):
    frames = np.stack([synth_image(20 + i, h, w) for i in range(2)])
    light = np.array([[0.31, 0.62, 0.71], [0.9, 0.5, 0.2]], np.float32)
    omega = np.float32(0.7)
    guard = 1e-6 if clip else 1e-10
    got = _np(prims.transmission_head_u8(frames, light, omega, guard, clip))
    for i in range(2):
        img = frames[i].astype(np.float32) / 255.0
        dark = np.min(img / (light[i].reshape(1, 1, 3) + guard), axis=2)            # SS:170 / ES:221 (float32)
        t0 = 1 - omega * dark
        if clip:
            t0 = np.clip(t0, 0.1, 1.0)
        assert t0.dtype == np.float32
        assert np.array_equal(got[i], t0), np.abs(got[i] - t0).max()


def test_quantise(cuda_device):
    rng = np.random.default_rng(0)
    for dt in (np.float32, np.float64):
        x = rng.random((37, 53, 3)).astype(dt)
        assert np.array_equal(_np(prims.quantise(x)), (x * 255).astype(np.uint8))


# ------------------------------------------------------------------ float-image primitives
def _rand_img(seed, h, w, dt, lo=0.0, hi=1.0):
    return (np.random.default_rng(seed).random((h, w, 3)) * (hi - lo) + lo).astype(dt)


@pytest.mark.parametrize("dt", [np.float32, np.float64])
@pytest.mark.parametrize("seed,h,w", [(1, 256, 256), (2, 97, 143)])
def test_atmospheric_light_float(cuda_device, dt, seed, h, w):
    """The descent of a NON byte-valued image: every level whose margin is above 1e-6 must agree (float64 sums
    here, dtype sums in NumPy); then the light is the same pixel."""
    img = _rand_img(seed, h, w, dt)
    img[h // 3: h // 3 + 20, w // 2: w // 2 + 30] *= 0.2          # structure, so that Q differs between quadrants
    trace = []
    want = o.quadtree_airlight(img, 1, trace)
    margins = [np.sort(np.asarray(t[-1]))[::-1] for t in trace]
    got = _np(prims.atmospheric_light(img))
    if all(m[0] - m[1] > 1e-6 for m in margins):
        assert np.array_equal(got, want), (got, want)
    assert got.dtype == dt


@pytest.mark.parametrize("r,eps", [(15, 1e-3), (20, 0.5), (7, 1e-2)])
def test_guided_filter_float(cuda_device, r, eps):
    rng = np.random.default_rng(4)
    guide, src = rng.random((120, 170)), rng.random((120, 170))
    got = _np(prims.guided_filter(guide, src, r, eps))
    assert np.abs(got - o.guided_filter(guide, src, r, eps)).max() < 1e-11


@pytest.mark.parametrize("dt", [np.float32, np.float64])
def test_transmission_recover_float(cuda_device, dt):
    img = _rand_img(5, 150, 210, dt, 0.15, 0.85)                   # example_usage.py:112
    light = o.quadtree_airlight(img, 1)
    plane = np.tile(light.reshape(1, 1, 3), (img.shape[0], img.shape[1], 1))
    t_ref = o.es_transmission(img, plane, omega=0.5, r=15)
    t = prims.transmission(img, light, 0.5, 15, 1e-3, 1e-10, False)
    assert np.abs(_np(t) - t_ref).max() < 1e-6
    rec = _np(prims.recover(img, t_ref, light, torch.float64))
    assert np.abs(rec - o.es_recover(img, t_ref, plane)).max() < 1e-12
    t_ss = o.ss_transmission(img.astype(np.float32), light.astype(np.float32), 0.3, 20, 0.5)
    t2 = prims.transmission(img.astype(np.float32), light.astype(np.float32), 0.3, 20, 0.5, 1e-6, True)
    assert np.abs(_np(t2) - t_ss).max() < 1e-6
    rec32 = _np(prims.recover(img.astype(np.float32), t_ss, light.astype(np.float32), torch.float32))
    assert rec32.dtype == np.float32
    assert np.abs(rec32 - o.ss_restore(img.astype(np.float32), light.astype(np.float32), t_ss)).max() < 1e-6


@pytest.mark.parametrize("dt,guard", [(np.float32, 1e-6), (np.float64, 1e-10)])
@pytest.mark.parametrize("lo,hi", [(10, 95), (2, 98), (0, 100)])
def test_percentile_stretch_float_exact(cuda_device, dt, guard, lo, hi):
    img = _rand_img(6, 131, 257, dt)
    img[:40, :60, 0] = 0.0                                          # ties at the ends, like a clipped recovered frame
    img[-30:, :, 2] = 1.0
    got = _np(prims.percentile_stretch(img, lo, hi, guard))
    want = o.percentile_stretch(img, lo, hi, guard)
    assert got.dtype == dt
    assert np.array_equal(got, want), np.abs(got - want).max()


def test_percentile_stretch_rejects_out_of_range(cuda_device):
    from underwater_image_enhancement_b200 import _lib
    img = _rand_img(7, 32, 32, np.float64) * 1.5
    with pytest.raises(_lib.UwieError):
        prims.percentile_stretch(img, 10, 90, 1e-10)


def test_gamma_scale_means(cuda_device):
    img = _rand_img(8, 64, 80, np.float64)
    assert np.abs(_np(prims.gamma(img, 1.0 / 1.2, True)) - o.es_gamma(img, 1.2)).max() < 1e-15
    img32 = img.astype(np.float32)
    assert np.abs(_np(prims.gamma(img32, 1.5, False)) - np.power(img32, 1.5)).max() < 1e-7
    want = o.ss_correct_cast(img32, "greenish")
    assert np.array_equal(_np(prims.scale_clip(img32, (1.0, 0.85, 1.0))), want)
    assert np.allclose(_np(prims.channel_means(img)), img.mean(axis=(0, 1)), rtol=1e-12)
