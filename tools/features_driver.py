"""features79 on 32 x 1080p frames (for an ncu launch list): python tools/features_driver.py [reps]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from underwater_image_enhancement_b200 import features, quality, vgg          # noqa: E402
from underwater_image_enhancement_b200.synth import synth_image               # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
base = np.stack([synth_image(60 + i, 1080, 1920) for i in range(8)])
frames = torch.from_numpy(np.concatenate([base] * 4)).cuda()
for _ in range(reps):
    features.extract_all_features_batch(frames)
    quality.quality_scores_batch(frames)
    p = torch.rand((4, 32), device="cuda") * torch.tensor([[10.0], [10.0], [0.5], [0.5]], device="cuda") + torch.tensor([[2.0], [85.0], [0.3], [0.9]], device="cuda")
    vgg.enhance_frames_u8(frames, p, True)
torch.cuda.synchronize()
print("ok")
