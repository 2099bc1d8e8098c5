"""Guided-filter geometry sweep: box sizes around the 5-column blocks of the marching kernels (k_guided.cu),
frames narrower than one strip, frames smaller than the box, odd widths - against the oracle's
cv2.boxFilter restatement (enhancement_strategies.py:209-249 through es_strategy)."""
from __future__ import annotations

import numpy as np
import pytest

from oracle import restatement as o
from underwater_image_enhancement_b200 import compat
from underwater_image_enhancement_b200.synth import synth_image

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("shape", [(97, 203), (61, 45), (200, 333)])
@pytest.mark.parametrize("radius", [1, 2, 4, 5, 6, 9, 10, 11, 33, 64])
def test_dehaze_radius_sweep(shape, radius):
    frame = synth_image(700 + radius, *shape)
    img = frame.astype(np.float32) / 255.0
    params = {"omega": 0.55, "guided_radius": radius, "L_low": 8, "L_high": 93}
    got = (compat.es_apply_strategy(frame, "medium_dehazing", params) * 255).astype(np.uint8)
    want = (o.es_strategy(img, "medium_dehazing", params) * 255).astype(np.uint8)
    d = np.abs(got.astype(np.int16) - want.astype(np.int16))
    assert d.max() <= 1                                             # north_star: max-abs 1/255 per pixel
    assert np.count_nonzero(d) <= max(2, 1e-3 * d.size)


def teardown_module():
    compat.close_engines()


@pytest.mark.parametrize("eps", [1e-12, 1e-3, 0.5])                 # 1e-12: the kernel variant with the IEEE routines
@pytest.mark.parametrize("radius", [3, 15, 20])
def test_six_library_custom_dehaze(eps, radius):
    """float32 library (clipped transmission, colour cast first) with a custom dehaze strategy."""
    from underwater_image_enhancement_b200 import _lib
    from underwater_image_enhancement_b200.engine import Engine
    frame = synth_image(900 + radius, 131, 277)
    s = _lib.Strategy()
    s.dehaze, s.omega, s.radius, s.eps = 1, 0.6, radius, eps
    s.n_pre = 1; s.pre_lo[0] = 4.0; s.pre_hi[0] = 96.0
    eng = Engine("six", 131, 277, max_batch=1, strategies=[s])
    try:
        got = eng.run_host(frame[None], want_all=True)["out"][0, 0]
    finally:
        eng.close()
    img = o.to_unit_f32(frame)
    cur = o.ss_correct_cast(img, o.ss_image_type(img))
    want = o.quantise_u8(o.ss_run_recipe(cur, (("dehaze", 0.6, radius, eps), ("stretch", 4, 96))))
    d = np.abs(got.astype(np.int16) - want.astype(np.int16))
    assert d.max() <= 1
    assert np.count_nonzero(d) <= max(2, 1e-3 * d.size)
