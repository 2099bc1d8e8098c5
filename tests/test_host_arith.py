"""The integer arithmetic the kernels share with the host (csrc/uwie_color.h, uwie_hd.h) compared
with cv2 / NumPy on the CPU.  tests/_hostcheck.so is a TEST-ONLY host compile of those headers; the
product library has no CPU path.  The exhaustive 2^24 sweeps run strided here to keep the CPU suite
short (UWIE_EXHAUSTIVE=1 runs all 16.7 M colours)."""
from __future__ import annotations

import ctypes as C
import os

import cv2
import numpy as np
import pytest

from oracle import restatement as o

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def hc(native_lib):
    lib = C.CDLL(os.path.join(HERE, "_hostcheck.so"))
    lib.hc_canny_class.restype = C.c_int
    return lib


def _colours():
    if os.environ.get("UWIE_EXHAUSTIVE") == "1":
        v = np.arange(1 << 24, dtype=np.uint32)
    else:
        rng = np.random.default_rng(0)
        v = np.concatenate([np.arange(0, 1 << 24, 37, dtype=np.uint32),
                            rng.integers(0, 1 << 24, 200_000, dtype=np.uint32)])
    rgb = np.stack([(v >> 16) & 255, (v >> 8) & 255, v & 255], axis=1).astype(np.uint8)
    return np.ascontiguousarray(rgb)


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def test_gray_matches_cv2(hc):
    rgb = _colours()
    out = np.empty(len(rgb), np.uint8)
    hc.hc_gray(_p(rgb), C.c_longlong(len(rgb)), _p(out))
    ref = cv2.cvtColor(rgb[None], cv2.COLOR_RGB2GRAY)[0]
    assert np.array_equal(out, ref)


def test_saturation_matches_cv2(hc):
    rgb = _colours()
    out = np.empty(len(rgb), np.uint8)
    hc.hc_sat(_p(rgb), C.c_longlong(len(rgb)), _p(out))
    ref = cv2.cvtColor(rgb[None], cv2.COLOR_RGB2HSV)[0, :, 1]
    assert np.array_equal(out, ref)


def test_rgb2lab_matches_cv2(hc):
    rgb = _colours()
    out = np.empty_like(rgb)
    hc.hc_rgb2lab(_p(rgb), C.c_longlong(len(rgb)), _p(out))
    ref = cv2.cvtColor(rgb[None], cv2.COLOR_RGB2LAB)[0]
    assert np.array_equal(out, ref)


def test_lab2rgb_matches_cv2(hc):
    lab = _colours()
    out = np.empty_like(lab)
    hc.hc_lab2rgb(_p(lab), C.c_longlong(len(lab)), _p(out))
    ref = cv2.cvtColor(lab[None], cv2.COLOR_LAB2RGB)[0]
    assert np.array_equal(out, ref)


@pytest.mark.parametrize("clip_limit", [1.5, 2.0, 3.0, 3.5, 4.0])
@pytest.mark.parametrize("shape", [(90, 160), (64, 64), (135, 241)])
def test_clahe_single_tile_lut_matches_cv2(hc, clip_limit, shape):
    """A 1x1 tile grid makes cv2's CLAHE a pure LUT of the tile histogram."""
    rng = np.random.default_rng(int(clip_limit * 10) + shape[0])
    plane = np.clip(rng.normal(110, 35, shape), 0, 255).astype(np.uint8)
    ref = cv2.createCLAHE(clipLimit=clip_limit, tileGridSize=(1, 1)).apply(plane)
    hist = np.bincount(plane.ravel(), minlength=256).astype(np.uint32)
    area = plane.size
    clip = max(int(clip_limit * area / 256), 1)
    lut = np.empty(256, np.uint8)
    hc.hc_clahe_lut(_p(hist), clip, area, _p(lut))
    assert np.array_equal(lut[plane], ref)


def test_equalize_lut_matches_cv2(hc):
    rng = np.random.default_rng(3)
    for plane in (np.clip(rng.normal(90, 30, (120, 160)), 0, 255).astype(np.uint8),
                  rng.integers(40, 200, (67, 121)).astype(np.uint8),
                  np.full((32, 32), 77, np.uint8)):
        hist = np.bincount(plane.ravel(), minlength=256).astype(np.uint32)
        lut = np.empty(256, np.uint8)
        hc.hc_equalize_lut(_p(hist), C.c_longlong(plane.size), _p(lut))
        assert np.array_equal(lut[plane], cv2.equalizeHist(plane))


def _canny_host(hc, gray):
    """Canny(50,150) through the shared NMS classifier + a plain flood fill for the hysteresis."""
    g = gray.astype(np.int32)
    p = np.pad(g, 1, mode="edge")
    dx = (p[:-2, 2:] + 2 * p[1:-1, 2:] + p[2:, 2:]) - (p[:-2, :-2] + 2 * p[1:-1, :-2] + p[2:, :-2])
    dy = (p[2:, :-2] + 2 * p[2:, 1:-1] + p[2:, 2:]) - (p[:-2, :-2] + 2 * p[:-2, 1:-1] + p[:-2, 2:])
    m = np.pad(np.abs(dx) + np.abs(dy), 1)
    h, w = gray.shape
    cls = np.zeros((h, w), np.uint8)
    for y in range(h):
        for x in range(w):
            cls[y, x] = hc.hc_canny_class(int(dx[y, x]), int(dy[y, x]), int(m[y + 1, x + 1]),
                                          int(m[y + 1, x]), int(m[y + 1, x + 2]), int(m[y, x + 1]),
                                          int(m[y + 2, x + 1]), int(m[y, x]), int(m[y, x + 2]),
                                          int(m[y + 2, x]), int(m[y + 2, x + 2]))
    edge = cls == 2
    stack = list(zip(*np.nonzero(edge)))
    while stack:
        y, x = stack.pop()
        for yy in range(max(0, y - 1), min(h, y + 2)):
            for xx in range(max(0, x - 1), min(w, x + 2)):
                if cls[yy, xx] and not edge[yy, xx]:
                    edge[yy, xx] = True
                    stack.append((yy, xx))
    return edge


@pytest.mark.parametrize("shape", [(1, 7), (2, 3), (40, 56), (31, 33)])
def test_canny_classifier_matches_cv2(hc, shape):
    from underwater_image_enhancement_b200.synth import synth_image
    frame = synth_image(11, max(shape[0], 48), max(shape[1], 64))[:shape[0], :shape[1]]
    gray = cv2.cvtColor(np.ascontiguousarray(frame), cv2.COLOR_RGB2GRAY)
    gray = (gray.astype(np.int32) * 2 % 256).astype(np.uint8)     # more contrast -> more edges
    assert np.array_equal(_canny_host(hc, gray), cv2.Canny(gray, 50, 150) > 0)


@pytest.mark.parametrize("f32", [1, 0])
def test_percentile_planner_matches_numpy(hc, f32):
    for n in (7, 1000, 19200, 120 * 160, 250 * 333, 1280 * 720, 1920 * 1080, 3840 * 2160):
        for q in (2, 3, 5, 10, 15, 20, 85, 90, 92, 95, 97, 98):
            k, g = C.c_longlong(), C.c_double()
            hc.hc_percentile_plan(C.c_longlong(n), C.c_double(q), f32, C.byref(k), C.byref(g))
            rk, rg = o.np_percentile_plan(n, q, bool(f32))
            assert (k.value, g.value) == (rk, float(rg)), (n, q)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_percentile_plan_reproduces_np_percentile(dtype):
    """The (k, gamma) plan + NumPy's lerp must reproduce np.percentile itself, bit for bit."""
    rng = np.random.default_rng(5)
    x = rng.random(19200).astype(dtype)
    xs = np.sort(x)
    for q in (5, 15, 20, 85, 95, 98):
        k, g = o.np_percentile_plan(x.size, q, dtype is np.float32)
        a, b = xs[k], xs[min(k + 1, x.size - 1)]
        d = b - a
        got = b - d * (1 - g) if g >= 0.5 else a + d * g
        assert dtype(got) == np.percentile(x, q)
