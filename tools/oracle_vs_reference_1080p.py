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

# ---- the 79 features (feature_extraction.py:249-297) on two of the frames of tests/test_gpu_features.py at 1080p; the
# texture group runs the reference's call sites on the oracle's stand-ins for scikit-image (tests/golden/make_golden_fe.py)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
from make_golden import REF, install_skimage_standin  # noqa: E402
install_skimage_standin()
import skimage  # noqa: E402
skimage.feature.local_binary_pattern = lambda image, P, R, method="default": o.sk_local_binary_pattern_uniform(image, P, R)
skimage.feature.graycomatrix = lambda image, distances, angles, levels=None, symmetric=False, normed=False: o.sk_graycomatrix(image, levels)
skimage.feature.graycoprops = o.sk_graycoprops
sys.path.insert(0, refload.reference_dir())
import feature_extraction as fe  # noqa: E402
fe_rows = []
for s in (21, 22):
    frame = synth_image(s, 1080, 1920)
    f = np.asarray(fe.FeatureExtractor.extract_all_features(frame.astype(np.float32) / 255.0), np.float64)
    g = o.fe_all(frame)
    fe_rows.append({"seed": s, "identical_values": int((f == g).sum()), "of": 79,
                    "max_rel_diff": float((np.abs(f - g) / np.maximum(np.abs(f), 1e-12)).max())})
    print("fe", fe_rows[-1], flush=True)
path = os.path.join(ROOT, 'profiles', 'r02_oracle_vs_reference_1080p.json')
doc = json.load(open(path))
doc["features79"] = fe_rows
json.dump(doc, open(path, 'w'), indent=1)

# ---- the VGG-path pieces (vgg_16_UIE.py:32-128, 435-466; use_trained_model.py:39-46) at 1080p
import torch  # noqa: E402
import torchvision.transforms as T  # noqa: E402
import use_trained_model as utm  # noqa: E402
import vgg_16_UIE as vggref  # noqa: E402
enh = vggref.DifferentiableEnhancement().eval()
pred = object.__new__(utm.EnhancementPredictor)
pred.input_size = 224
pred.normalize = T.Normalize(mean=[0.485, 0.456, 0.406], std=[0.229, 0.224, 0.225])
vgg_rows = []
for s, lo, hi, om, ga in ((60, 4.0, 96.0, 0.55, 1.3), (61, 12.0, 88.0, 0.8, 0.9)):
    frame = synth_image(s, 1080, 1920)
    img = frame.astype(np.float32) / 255.0
    t = torch.from_numpy(img).permute(2, 0, 1).unsqueeze(0)
    def p(x):
        return torch.tensor([[float(x)]], dtype=torch.float32)
    with torch.no_grad():
        full = enh(t, {"L_low": p(lo), "L_high": p(hi), "omega": p(om), "gamma": p(ga)})[0].numpy()
        nog = enh(t, {"L_low": p(lo), "L_high": p(hi), "omega": p(om)})[0].numpy()
    x = np.ascontiguousarray(img.transpose(2, 0, 1))[None]
    o_full = o.vgg_enhance(x, [lo], [hi], [om], [ga])[0]
    o_nog = o.vgg_enhance(x, [lo], [hi], [om], None)[0]
    feats = vggref.extract_all_features(img)
    pre = pred._preprocess_for_vgg(img)[0].numpy()
    vgg_rows.append({"seed": s, "enhance_no_gamma_identical": bool(np.array_equal(nog, o_nog)),
                     "enhance_max_abs_diff": float(np.abs(full - o_full).max()),
                     "features_max_rel_diff": float((np.abs(feats - o.vgg_features(frame)) / np.maximum(np.abs(feats), 1e-12)).max()),
                     "preprocess_identical": bool(np.array_equal(pre, o.vgg_preprocess(frame, 224)))})
    print("vgg", vgg_rows[-1], flush=True)
doc = json.load(open(path))
doc["vgg_path"] = vgg_rows
json.dump(doc, open(path, 'w'), indent=1)
