"""BASELINE.json configs[2] and configs[3] as parity cases (configs[1] is the bench line, configs[0] the
CPU-runnable fixture size, configs[4] the sharded stream covered by tests/test_shard_gloo.py + bench --gpus N).

configs[2]: CLAHE and dark-channel dehaze ONLY on 3840x2160 frames — an engine with those two strategies.
configs[3]: the VGG-guided path at 1080p — enhancer, statistics and network input of a small batch."""
from __future__ import annotations

import numpy as np
import pytest
import torch

from oracle import restatement as o
from underwater_image_enhancement_b200 import compat, vgg
from underwater_image_enhancement_b200.synth import synth_image

pytestmark = pytest.mark.gpu


def test_config2_clahe_and_dehaze_only_at_2160p():
    frame = synth_image(50, 2160, 3840)
    img = frame.astype(np.float32) / 255.0
    for name, params in (("clahe_enhancement", {"clip_limit": 2.0, "tile_grid_size": (8, 8)}),
                         ("light_enhancement", {"omega": 0.4, "guided_radius": 10, "L_low": 15, "L_high": 95})):
        got = (compat.es_apply_strategy(frame, name, params) * 255).astype(np.uint8)
        want = (o.es_strategy(img, name, params) * 255).astype(np.uint8)
        d = np.abs(got.astype(np.int16) - want.astype(np.int16))
        assert d.max() <= 1, name                                   # north_star: max-abs 1/255 per pixel
        assert np.count_nonzero(d) <= 1e-3 * d.size, name
    compat.close_engines()


def test_config3_vgg_path_at_1080p():
    frames = np.stack([synth_image(60 + i, 1080, 1920) for i in range(2)])
    img = torch.from_numpy(frames.astype(np.float32) / 255.0).permute(0, 3, 1, 2).contiguous().cuda()
    lo, hi, om, ga = [10.0, 4.0], [90.0, 97.0], [0.6, 0.8], [1.2, 1.0]
    got = vgg.DifferentiableEnhancement()(img, {"L_low": lo, "L_high": hi, "omega": om, "gamma": ga}).cpu().numpy()
    want = o.vgg_enhance(img.cpu().numpy(), lo, hi, om, ga)
    assert np.abs(got - want).max() <= 1e-6
    feats = vgg.extract_all_features(frames).cpu().numpy()
    pre = vgg.preprocess_for_vgg(frames).cpu().numpy()
    for i in range(2):
        assert np.allclose(feats[i], o.vgg_features(frames[i]), rtol=1e-5, atol=1e-7)
        assert np.array_equal(pre[i], o.vgg_preprocess(frames[i]))
