"""NumPy mirror of the index arithmetic of csrc/k_dct_fft.cu (Stockham autosort FFT with radices 4, 2, 3, 5; Makhoul's
N-point DCT-II; two real sequences per complex FFT), checked against np.fft / cv2.dct / scipy on the CPU.  The CUDA file
follows these loops line by line; the device result itself is checked against the oracle in tests/test_gpu_features.py."""
from __future__ import annotations

import numpy as np
import pytest


def factors(n):
    f = []
    for r in (4, 2, 3, 5):
        while n % r == 0:
            f.append(r)
            n //= r
    return f if n == 1 else None


def stockham(x, fac):
    N = len(x)
    W = np.exp(-2j * np.pi * np.arange(N) / N).astype(np.complex64)
    x = x.astype(np.complex64).copy()
    y = np.empty_like(x)
    n, s = N, 1
    for r in fac:
        m = n // r
        wr = np.exp(-2j * np.pi * np.arange(r) / r)
        for t in range(N // r):                           # one butterfly per (p, q), as a thread takes it
            p, q = divmod(t, s)
            a = [x[q + s * (p + m * j)] for j in range(r)]
            for k in range(r):
                assert p * k * s < N                      # the twiddle index needs no reduction
                y[q + s * (r * p + k)] = sum(a[j] * wr[(j * k) % r] for j in range(r)) * W[p * k * s]
        x, y = y, x
        n, s = m, s * r
    return x


def dct_pair(x1, x2):
    N = len(x1)

    def reorder(x):
        v = np.empty(N, np.float32)
        c = np.arange(N)
        v[np.where(c & 1, N - 1 - (c >> 1), c >> 1)] = x
        return v

    Z = stockham(reorder(x1) + 1j * reorder(x2), factors(N))
    k = np.arange(N)
    Zr = np.conj(Z[(N - k) % N])
    V1, V2 = (Z + Zr) / 2, (Z - Zr) / 2j
    P = np.exp(-1j * np.pi * k / (2 * N)) * np.where(k == 0, np.sqrt(1 / N), np.sqrt(2 / N))
    return (P * V1).real, (P * V2).real


@pytest.mark.parametrize("n", [8, 12, 30, 45, 60, 120, 135])
def test_mixed_radix_fft_and_dct(n):
    from scipy.fft import dct
    rng = np.random.default_rng(n)
    x1, x2 = rng.random(n).astype(np.float32), rng.random(n).astype(np.float32)
    Z = stockham(x1 + 1j * x2, factors(n))
    assert np.abs(Z - np.fft.fft(x1.astype(np.float64) + 1j * x2)).max() < 4e-6 * n ** 0.5
    a, b = dct_pair(x1, x2)
    assert np.abs(a - dct(x1.astype(np.float64), norm="ortho")).max() < 2e-6
    assert np.abs(b - dct(x2.astype(np.float64), norm="ortho")).max() < 2e-6
    if n % 2 == 0:
        import cv2
        assert np.abs(a - cv2.dct(x1.reshape(1, -1)).ravel()).max() < 2e-6


def test_sizes_of_the_baseline_configs_factor():
    for n in (1920, 1080, 3840, 2160, 1280, 720):
        assert factors(n) is not None and len(factors(n)) <= 14
    assert factors(33) is None and factors(97) is None


@pytest.mark.parametrize("h,w", [(12, 30), (45, 8), (9, 25)])
def test_two_passes_give_cv2_dct_and_the_five_sums(h, w):
    """Rows in pairs, then columns in pairs (a lone last row / column rides with zeros), as the two kernels do it; the five
    sums of feature_extraction.py:142-156 from the coefficients."""
    from scipy.fft import dctn
    rng = np.random.default_rng(h * 100 + w)
    g = rng.random((h, w)).astype(np.float32)
    T = np.empty((h, w), np.float32)
    for r in range(0, h, 2):
        a, b = dct_pair(g[r], g[r + 1] if r + 1 < h else np.zeros(w, np.float32))
        T[r] = a
        if r + 1 < h:
            T[r + 1] = b
    Z = np.empty((h, w), np.float32)
    for c in range(0, w, 2):
        a, b = dct_pair(T[:, c], T[:, c + 1] if c + 1 < w else np.zeros(h, np.float32))
        Z[:, c] = a
        if c + 1 < w:
            Z[:, c + 1] = b
    ref = dctn(g.astype(np.float64), norm="ortho")
    assert np.abs(Z - ref).max() < 4e-6
    e, er = Z.astype(np.float64) ** 2, ref ** 2
    for sl in ((slice(0, h // 4), slice(0, w // 4)), (slice(h // 4, h // 2), slice(w // 4, w // 2)), (slice(h // 2, h), slice(w // 2, w))):
        assert abs(e[sl].sum() - er[sl].sum()) <= 1e-5 * max(er[sl].sum(), 1e-9) + 1e-9
    assert abs(e.sum() - er.sum()) <= 1e-5 * er.sum()
