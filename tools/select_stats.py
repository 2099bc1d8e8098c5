"""How many percentile targets the candidate lists resolved and how many fell back to the streaming
levels B / C, on the bench's synthetic 1080p frames (GPU box).   python tools/select_stats.py [--lib six|es]"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from underwater_image_enhancement_b200.engine import Engine            # noqa: E402
from underwater_image_enhancement_b200.synth import synth_image        # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--lib", default="six")
ap.add_argument("--frames", type=int, default=8)
args = ap.parse_args()
h, w = 1080, 1920
eng = Engine(args.lib, h, w, max_batch=args.frames)
fr = torch.from_numpy(np.stack([synth_image(s, h, w) for s in range(args.frames)])).cuda()
eng.run(fr)
eng.check()
st = eng.debug("sel_stats", np.uint32)
print(f"select_stats lib={args.lib} frames={args.frames} resolved_from_lists={int(st[0])} streamed={int(st[1])}")
