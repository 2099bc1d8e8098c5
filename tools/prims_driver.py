"""One pass of the two isolated configs[2] kernels at 3840x2160 x 32 (for ncu): python tools/prims_driver.py [reps]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from underwater_image_enhancement_b200 import prims                         # noqa: E402
from underwater_image_enhancement_b200.synth import synth_image            # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
import cv2                                                                  # noqa: E402
h, w, batch = 2160, 3840, 32
base = np.stack([synth_image(50 + i, h, w) for i in range(4)])
lplanes = np.stack([cv2.cvtColor(f, cv2.COLOR_RGB2LAB)[:, :, 0] for f in base])
planes = torch.from_numpy(np.concatenate([lplanes] * (batch // 4))).cuda()
frames = torch.from_numpy(np.concatenate([base] * (batch // 4))).cuda()
light = torch.tensor([[0.31, 0.62, 0.71]] * batch, dtype=torch.float32, device="cuda")
pout, tout = torch.empty_like(planes), torch.empty((batch, h, w), dtype=torch.float32, device="cuda")
for _ in range(reps):
    prims.clahe_lab(planes, 2.0, (8, 8), out=pout)
    prims.transmission_head_u8(frames, light, 0.5, 1e-6, True, out=tout)
torch.cuda.synchronize()
print("ok")
