import sys, time, json
import os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from oracle import refload, restatement as o
from underwater_image_enhancement_b200.synth import synth_image
mods = refload.load_reference()
rep = {}
for lib, seeds in (("six", (0, 1, 2)), ("es", (0, 1))):
    rows = []
    for s in seeds:
        frame = synth_image(s, 1080, 1920)
        t0 = time.time()
        ref = refload.p6_frame(mods, frame) if lib == "six" else refload.p5_frame(mods, frame)
        t1 = time.time()
        ora = o.p6_frame(frame) if lib == "six" else o.p5_frame(frame)
        t2 = time.time()
        diff = [int(np.count_nonzero(ref["out"][k] != ora["out"][k])) for k in range(len(ref["out"]))]
        rel = float((np.abs(ref["scores"] - ora["scores"]) / np.maximum(np.abs(ref["scores"]), 1e-9)).max())
        rows.append({"seed": s, "bytes_differing_per_strategy": diff, "max_score_rel_diff": rel,
                     "reference_s": round(t1 - t0, 1), "oracle_s": round(t2 - t1, 1)})
        print(lib, rows[-1], flush=True)
    rep[lib] = rows
json.dump({"what": "oracle/restatement.py against the unmodified reference (oracle/refload.py) at 1920x1080, CPU", "frames": rep},
          open(os.path.join(ROOT, 'profiles', 'r02_oracle_vs_reference_1080p.json'), 'w'), indent=1)
