"""The reference-named entry points (compat.py) against the fixtures of the unmodified reference."""
from __future__ import annotations

import numpy as np
import pytest

from conftest import golden_paths, load_golden
from underwater_image_enhancement_b200 import compat
from underwater_image_enhancement_b200.engine import ES_NAMES, SS_NAMES

pytestmark = pytest.mark.gpu

CFG = {   # config.py:28-75
    "strong_dehazing": {"omega": 0.5, "guided_radius": 15, "L_low": 10, "L_high": 95, "gamma": 1.2, "apply_gamma": True},
    "medium_dehazing": {"omega": 0.6, "guided_radius": 20, "L_low": 15, "L_high": 92, "apply_gamma": True},
    "light_enhancement": {"omega": 0.4, "guided_radius": 10, "L_low": 15, "L_high": 95, "apply_gamma": False},
    "clahe_enhancement": {"clip_limit": 2.0, "tile_grid_size": (8, 8), "apply_gamma": False},
    "histogram_equalization": {"L_low": 10, "L_high": 95},
}
WEIGHTS = {"contrast": 0.25, "sharpness": 0.20, "entropy": 0.15, "saturation": 0.15, "brightness": 0.15,
           "edge_density": 0.10}


@pytest.fixture(scope="module")
def g():
    return load_golden(golden_paths()[0])


def test_apply_strategy_matches_reference(g):
    img = g["frame"].astype(np.float32) / 255.0                                  # main.py:108
    for i, name in enumerate(ES_NAMES):
        out = compat.es_apply_strategy(img, name, CFG[name])
        assert out.dtype == np.float64 and out.shape == img.shape
        assert np.array_equal((out * 255).astype(np.uint8), g["p5_out"][i]), name
    with pytest.raises(ValueError):
        compat.es_apply_strategy(img, "white_balance", {})                      # ES:500-501
    rnd = np.random.default_rng(0).random((64, 64, 3))                          # not made from bytes: the primitive route
    out = compat.es_apply_strategy(rnd, "clahe_enhancement", {})                # example_usage.py:27-31
    assert out is not rnd and out.shape == rnd.shape and out.dtype == np.float64
    bad = np.zeros((16, 16))                                                    # a failure inside: prints, returns input
    assert compat.es_apply_strategy(bad, "clahe_enhancement", {}) is bad        # ES:506-508


def test_enhance_and_select_matches_reference(g):
    img = g["frame"].astype(np.float32) / 255.0
    best, best_img, totals, subs = compat.enhance_and_select(img, CFG, WEIGHTS)
    want = int(np.argmax(g["p5_scores"][:, 0]))
    assert best == ES_NAMES[want]
    assert np.array_equal((best_img * 255).astype(np.uint8), g["p5_out"][want])
    for i, name in enumerate(ES_NAMES):
        assert abs(totals[name] - g["p5_scores"][i, 0]) <= 1e-4 * abs(g["p5_scores"][i, 0])
    assert compat.enhance_and_select(g["frame"])[0] == best                      # uint8 input, presets


def test_six_strategies_and_file_loop(g, tmp_path):
    import cv2
    kind, outs, scores = compat.six_strategies(g["frame"], with_scores=True)
    assert kind == ("normal", "greenish", "bluish")[int(g["p6_type"])]
    assert compat.detect_image_type(g["frame"].astype(np.float32) / 255.0) == kind
    for i, name in enumerate(SS_NAMES):
        assert outs[name].dtype == np.float32
        assert np.array_equal((outs[name] * 255).astype(np.uint8), g["p6_out"][i]), name
        assert np.allclose(scores[name], g["p6_scores"][i], rtol=1e-4)
    src = tmp_path / "in"; src.mkdir()
    cv2.imwrite(str(src / "a.png"), cv2.cvtColor(g["frame"], cv2.COLOR_RGB2BGR))
    cv2.imwrite(str(src / "b.png"), cv2.cvtColor(g["frame"][::-1].copy(), cv2.COLOR_RGB2BGR))
    (src / "notes.txt").write_text("x")
    log = compat.process_all_images_all_strategies(src, tmp_path / "out")
    assert len(log) == 12 and all(r["status"] == "success" for r in log)
    for i, name in enumerate(SS_NAMES):
        back = cv2.cvtColor(cv2.imread(str(tmp_path / "out" / f"a_{name}.png")), cv2.COLOR_BGR2RGB)
        assert np.array_equal(back, g["p6_out"][i])
    compat.close_engines()
