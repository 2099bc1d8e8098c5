"""Small-shape pass over every entry point of libuwie_b200.so, for compute-sanitizer (one tool per gpurun call):

    compute-sanitizer --tool memcheck  python tools/sanitize_driver.py
    compute-sanitizer --tool racecheck python tools/sanitize_driver.py
    compute-sanitizer --tool initcheck python tools/sanitize_driver.py

Shapes are tiny so that the 10-100x slow-down of the tools stays within a minute or two; they still cross CLAHE tile
borders, several Canny tiles (rim passes) and ~7 quadtree levels.
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from underwater_image_enhancement_b200 import features, prims, quality, vgg      # noqa: E402
from underwater_image_enhancement_b200.engine import Engine                       # noqa: E402
from underwater_image_enhancement_b200.synth import synth_image                   # noqa: E402

h, w = 100, 150
frames = np.stack([synth_image(s, h, w) for s in (0, 1, 2)])
for lib in ("six", "es"):
    eng = Engine(lib, h, w, max_batch=2)
    r = eng.run(torch.from_numpy(frames[:2]).cuda(), want_all=True, want_best=True)
    eng.check()
    rh = eng.run_host(frames, want_all=True, want_best=True)          # 3 frames in chunks of 2: both lanes, both slots
    print(lib, "best", r["best"].cpu().tolist(), rh["best"].tolist(), "launches", eng.last_launches)
    eng.close()
d = torch.from_numpy(frames).cuda()
print("qa", quality.quality_scores_batch(d)[:, 0].cpu().tolist())
print("fe", features.extract_all_features_batch(d)[:, :3].cpu().tolist())
img = (d.permute(0, 3, 1, 2).float() / 255.0).contiguous()
p = torch.tensor([[5.0, 10.0, 2.0], [95.0, 90.0, 98.0], [0.5, 0.7, 0.4], [1.2, 0.9, 1.4]], device="cuda")
a, _ = vgg.DifferentiableEnhancement().forward_dev(img, p, True)
b, _ = vgg.enhance_frames_u8(d, p, True)
print("vgg", float((a - b).abs().max()), vgg.extract_all_features(d).shape, vgg.preprocess_for_vgg(d, 64).shape)
import cv2                                                                         # noqa: E402
planes = np.stack([cv2.cvtColor(f, cv2.COLOR_RGB2LAB)[:, :, 0] for f in frames])
print("clahe", int(prims.clahe_lab(planes, 2.0, (8, 8)).sum()), int(prims.clahe_lab(np.ascontiguousarray(np.tile(planes, (1, 2, 2))[:, :176, :256]), 3.0, (4, 4)).sum()))
print("eq", int(prims.equalize_hist(frames).sum()))
light = torch.tensor([[0.3, 0.6, 0.7]] * 3, device="cuda")
print("thead", float(prims.transmission_head_u8(frames, light, 0.5, 1e-6, True).sum()))
rng = np.random.default_rng(0)
for dt in (np.float32, np.float64):
    x = rng.random((h, w, 3)).astype(dt)
    A = prims.atmospheric_light(x)
    t = prims.transmission(x, A, 0.5, 15, 1e-3, 1e-10, False)
    rec = prims.recover(x, t, A, torch.float64 if dt == np.float64 else torch.float32)
    st = prims.percentile_stretch(rec, 10, 95, 1e-10)
    g = prims.gamma(st, 1 / 1.2, True)
    print("float", dt.__name__, A.cpu().tolist(), float(g.sum()), prims.region_score(x), prims.brightest_pixel(x).cpu().tolist(),
          prims.channel_means(x).cpu().tolist(), float(prims.scale_clip(x, (1, .85, 1)).sum()), int(prims.quantise(x).sum()))
torch.cuda.synchronize()
print("sanitize_driver: done")
