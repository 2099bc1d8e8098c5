"""Generate tests/golden/*.npz by running the UNMODIFIED reference in the build container.

    python tests/golden/make_golden.py            # needs /root/reference (read-only mount)

The reference is pure Python and imports fine here once a module named `skimage` exists
(scikit-image is neither installed nor installable: no network).  The only skimage function the
hot path calls is skimage.measure.shannon_entropy (quality_assessment.py:70), which in
scikit-image >= 0.19 is `scipy.stats.entropy(np.unique(x, return_counts=True)[1], base=2)`; the
stand-in below is exactly that.  matplotlib/seaborn are plot-only and are not imported by the
modules used here.

The fixtures pin oracle/restatement.py (tests/test_oracle_golden.py) and, through it, the CUDA
path.  /root/reference does not exist on the GPU box, so nothing at test time imports it.
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("UWIE_REF_DIR", "/root/reference")
sys.path.insert(0, ROOT)


def install_skimage_standin():
    from scipy.stats import entropy as scipy_entropy

    def shannon_entropy(image, base=2):
        _, counts = np.unique(image, return_counts=True)
        return scipy_entropy(counts, base=base)

    sk = types.ModuleType("skimage")
    for sub in ("color", "exposure", "measure", "feature", "filters"):
        m = types.ModuleType("skimage." + sub)
        setattr(sk, sub, m)
        sys.modules["skimage." + sub] = m
    sk.measure.shannon_entropy = shannon_entropy
    sys.modules["skimage"] = sk


def load_reference():
    install_skimage_standin()
    sys.path.insert(0, REF)
    import config as ref_config                     # noqa: E402
    import enhancement_strategies as ref_es         # noqa: E402
    import quality_assessment as ref_qa             # noqa: E402
    import six_stadigy as ref_ss                    # noqa: E402
    return ref_config, ref_es, ref_qa, ref_ss


QA_KEYS = ("contrast", "sharpness", "entropy", "saturation", "brightness", "edge_density",
           "colorfulness", "naturalness")

# (seed, H, W): small enough to keep the fixtures to a few MB, large enough for 8x8 CLAHE tiles,
# r=20 guided-filter windows and ~7 quadtree levels; one size is not divisible by 8 (CLAHE pad).
CASES = ((0, 120, 160), (1, 120, 160), (2, 120, 160), (3, 135, 241), (4, 96, 128), (7, 250, 333))


def main():
    from underwater_image_enhancement_b200.synth import synth_image
    cfg, es, qa, ss = load_reference()
    ss_funcs = (ss.EnhancementStrategies.strategy1_strong_dehazing,
                ss.EnhancementStrategies.strategy2_medium_dehazing,
                ss.EnhancementStrategies.strategy3_light_dehazing,
                ss.EnhancementStrategies.strategy4_clahe_enhancement,
                ss.EnhancementStrategies.strategy5_white_balance,
                ss.EnhancementStrategies.strategy6_histogram_eq)
    for seed, h, w in CASES:
        frame = synth_image(seed, h, w)
        img = frame.astype(np.float32) / 255.0                      # SS:406 / main.py:108

        # ---- P6: six_stadigy.py:409-431 + comprehensive_assessment per output
        kind = ss.detect_image_type(img)
        cur = ss.color_correction(img, kind)
        light = np.array(ss.EnhancementStrategies.estimate_atmospheric_light(cur, min_size=1))
        p6_out, p6_f, p6_scores = [], [], []
        for fn in ss_funcs:
            enh = fn(cur)
            p6_f.append(enh.astype(np.float32))
            p6_out.append((enh * 255).astype(np.uint8))
            total, sub = qa.QualityAssessment.comprehensive_assessment(enh)
            p6_scores.append([total] + [sub[k] for k in QA_KEYS])

        # ---- P5: main.py:122-146 (Config.STRATEGIES, Config.QUALITY_WEIGHTS)
        p5_out, p5_f, p5_scores = [], [], []
        for key, params in cfg.Config.STRATEGIES.items():
            enh = es.EnhancementStrategies.apply_strategy(img, key, params)
            p5_f.append(np.asarray(enh, dtype=np.float64))
            p5_out.append((enh * 255).astype(np.uint8))
            total, sub = qa.QualityAssessment.comprehensive_assessment(
                enh, weights=cfg.Config.QUALITY_WEIGHTS)
            p5_scores.append([total] + [sub[k] for k in QA_KEYS])
        plane = es.EnhancementStrategies.estimate_atmospheric_light(img, min_size=1)

        path = os.path.join(HERE, f"ref_seed{seed}_{h}x{w}.npz")
        extra = {}
        if h * w <= 120 * 160:   # float planes only for the small cases (fixture size)
            extra = dict(p6_s3_float=p6_f[2],      # the one SS output that is not uint8-valued
                         p5_strong_float=p5_f[0])
        np.savez_compressed(
            path,
            frame=frame,
            p6_type=np.array(["normal", "greenish", "bluish"].index(kind)),
            p6_light=light.astype(np.float32),
            p6_out=np.stack(p6_out),
            p6_scores=np.asarray(p6_scores, dtype=np.float64),
            p5_light=np.asarray(plane[0, 0, :], dtype=np.float32),
            p5_out=np.stack(p5_out),
            p5_scores=np.asarray(p5_scores, dtype=np.float64),
            **extra,
        )
        print(path, os.path.getsize(path) // 1024, "KiB", "type", kind,
              "p6 totals", np.round(np.asarray(p6_scores)[:, 0], 3))


if __name__ == "__main__":
    main()
