import numpy as np, torch, json, sys, os
sys.path.insert(0, os.getcwd())
from oracle import restatement as o
from underwater_image_enhancement_b200.engine import Engine
from underwater_image_enhancement_b200.synth import synth_image
H, W = 1080, 1920
seeds = (0, 1, 2)
frames = np.stack([synth_image(s, H, W) for s in seeds])
eng = Engine("six", H, W, max_batch=3)
res = eng.run(torch.from_numpy(frames).cuda(), want_all=True, want_best=True)
eng.check()
rep = []
for i, s in enumerate(seeds):
    ref = o.p6_frame(frames[i])
    out = res["out"][i].cpu().numpy()
    row = {"seed": s, "flips": [], "max_d": [], "best_same": int(res["best"][i]) == ref["best"]}
    for k in range(6):
        d = np.abs(out[k].astype(np.int16) - ref["out"][k].astype(np.int16))
        row["flips"].append(int(np.count_nonzero(d))); row["max_d"].append(int(d.max()))
    sc = res["scores"][i].cpu().numpy()
    row["score_rel"] = float((np.abs(sc - ref["scores"]) / np.maximum(np.abs(ref["scores"]), 1e-9)).max())
    rep.append(row); print(row)
json.dump({"env": os.environ.get("UWIE_GF_AB32"), "bytes_per_strategy": H * W * 3, "frames": rep}, open("gpurun_out/ab32_flips.json", "w"), indent=1)
