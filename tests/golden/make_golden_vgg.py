"""Generate tests/golden/vgg_*.npz by running the UNMODIFIED reference's VGG-path pieces here.

    python tests/golden/make_golden_vgg.py        # needs /root/reference (read-only mount)

Pinned: DifferentiableEnhancement.forward (vgg_16_UIE.py:32-128, torch CPU), extract_all_features
(vgg_16_UIE.py:435-466) and EnhancementPredictor._preprocess_for_vgg (use_trained_model.py:39-46).
The predictor object is built without its checkpoint-loading __init__ (only `input_size` and
`normalize` are used by the method pinned here).
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

from make_golden import REF, install_skimage_standin  # noqa: E402

# (seed, H, W, L_low, L_high, omega, gamma)
CASES = ((0, 120, 160, 10.0, 90.0, 0.6, 1.2), (3, 135, 241, 3.5, 97.25, 0.85, 1.45),
         (7, 250, 333, 14.0, 66.0, 0.33, 1.0), (9, 448, 448, 5.0, 95.0, 0.5, 1.1))


def main():
    from underwater_image_enhancement_b200.synth import synth_image
    install_skimage_standin()
    sys.path.insert(0, REF)
    import torchvision.transforms as T
    import use_trained_model as utm
    import vgg_16_UIE as vgg
    enh = vgg.DifferentiableEnhancement().eval()
    pred = object.__new__(utm.EnhancementPredictor)
    pred.input_size = 224
    pred.normalize = T.Normalize(mean=[0.485, 0.456, 0.406], std=[0.229, 0.224, 0.225])
    for seed, h, w, lo, hi, om, ga in CASES:
        frame = synth_image(seed, h, w)
        img = frame.astype(np.float32) / 255.0                                    # UTM:120
        t = torch.from_numpy(img).permute(2, 0, 1).unsqueeze(0)                   # UTM:48-51
        def p(x):
            return torch.tensor([[float(x)]], dtype=torch.float32)
        with torch.no_grad():
            full = enh(t, {"L_low": p(lo), "L_high": p(hi), "omega": p(om), "gamma": p(ga)})[0].numpy()
            stretch_only = enh(t, {"L_low": p(lo), "L_high": p(hi)})[0].numpy()
        feats = vgg.extract_all_features(img)
        pre = pred._preprocess_for_vgg(img)[0].numpy()
        path = os.path.join(HERE, f"vgg_seed{seed}_{h}x{w}.npz")
        np.savez_compressed(path, frame=frame, params=np.array([lo, hi, om, ga], np.float32),
                            enhanced=full.astype(np.float32), stretched=stretch_only.astype(np.float32),
                            features=feats.astype(np.float32), preprocessed=pre.astype(np.float32))
        print(path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
