"""Seeded synthetic underwater-tinted frames.

There is no dataset on the build or GPU boxes, so every test and bench input comes from
this generator.  It is pure NumPy on the host so that the CPU oracle and the CUDA path see
byte-identical uint8 frames.

frame(seed) = trunc(255 * tint * clip(low(x, y) + discs + a * (U[0,1)^3 - 0.5), 0, 1))

* low   : smooth illumination field, 0.45 + 0.2 sin(x/97 + seed) + 0.2 cos(y/61 - seed)
* discs : 40 hard-edged discs with a per-disc RGB offset -> real Canny contours
* a     : noise amplitude, cycled over (0.02, 0.04, 0.08)
* tint  : seed % 3 == 0 bluish water, 1 greenish water, 2 neutral, so that every branch of
          the colour-cast detector (six_stadigy.py:292-302 in the reference) fires.

Pure uniform noise must not be used as a benchmark input: it saturates the edge-density and
colourfulness scores and hides Canny / colour errors.
"""
from __future__ import annotations

import numpy as np

TINTS = (
    (0.35, 0.75, 0.85),  # bluish
    (0.35, 0.85, 0.60),  # greenish
    (0.80, 0.80, 0.80),  # neutral
)
NOISE = (0.02, 0.04, 0.08)
N_DISCS = 40


def synth_image(seed: int, height: int = 1080, width: int = 1920) -> np.ndarray:
    """Return one uint8 RGB frame of shape (height, width, 3)."""
    rng = np.random.default_rng(int(seed))
    x = np.arange(width, dtype=np.float32)
    y = np.arange(height, dtype=np.float32)
    s = np.float32(seed % 1000)
    low = (np.float32(0.45)
           + np.float32(0.2) * np.sin(x / np.float32(97.0) + s)[None, :]
           + np.float32(0.2) * np.cos(y / np.float32(61.0) - s)[:, None]).astype(np.float32)
    img = np.repeat(low[:, :, None], 3, axis=2)

    cx = rng.uniform(0, width, N_DISCS)
    cy = rng.uniform(0, height, N_DISCS)
    rad = rng.uniform(0.01, 0.06, N_DISCS) * width
    off = rng.uniform(-0.25, 0.25, (N_DISCS, 3)).astype(np.float32)
    for i in range(N_DISCS):
        x0, x1 = int(max(0, cx[i] - rad[i])), int(min(width, cx[i] + rad[i] + 1))
        y0, y1 = int(max(0, cy[i] - rad[i])), int(min(height, cy[i] + rad[i] + 1))
        if x0 >= x1 or y0 >= y1:
            continue
        yy, xx = np.mgrid[y0:y1, x0:x1]
        m = (xx - cx[i]) ** 2 + (yy - cy[i]) ** 2 <= rad[i] ** 2
        img[y0:y1, x0:x1][m] += off[i]

    amp = np.float32(NOISE[(seed // 3) % 3])
    noise = rng.random((height, width, 3), dtype=np.float32)
    img += amp * (noise - np.float32(0.5))
    np.clip(img, 0.0, 1.0, out=img)
    img *= np.asarray(TINTS[seed % 3], dtype=np.float32)
    img *= np.float32(255.0)
    return img.astype(np.uint8)


def synth_batch(first_seed: int, count: int, height: int = 1080, width: int = 1920,
                out: np.ndarray | None = None) -> np.ndarray:
    """uint8 (count, height, width, 3); frame i uses seed first_seed + i."""
    if out is None:
        out = np.empty((count, height, width, 3), dtype=np.uint8)
    for i in range(count):
        out[i] = synth_image(first_seed + i, height, width)
    return out
