"""TEST / BENCH INFRASTRUCTURE ONLY — loader of the UNMODIFIED reference (pure Python) for the CPU arm.

The reference is a directory of scripts (no setup.py / pyproject.toml, so `pip install --target` has
nothing to install).  In the build container it lives read-only at /root/reference; `__graft_entry__.build()`
copies its *.py files verbatim into the git-ignored `baseline/_ref/` so that they travel to the GPU box with
gpurun (they never enter the history).  Nothing under underwater_image_enhancement_b200/ imports this file.

The only third-party module the path needs and this image lacks is scikit-image; the single function the
path calls from it is skimage.measure.shannon_entropy (quality_assessment.py:70), which in scikit-image
>= 0.19 is scipy.stats.entropy(np.unique(x, return_counts=True)[1], base=2): the stand-in below is that.
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CANDIDATES = (os.environ.get("UWIE_REF_DIR", ""), os.path.join(ROOT, "baseline", "_ref"), "/root/reference")
QA_KEYS = ("contrast", "sharpness", "entropy", "saturation", "brightness", "edge_density",
           "colorfulness", "naturalness")


def reference_dir() -> str | None:
    for d in CANDIDATES:
        if d and os.path.exists(os.path.join(d, "six_stadigy.py")):
            return d
    return None


def install_skimage_standin() -> None:
    if "skimage" in sys.modules and not getattr(sys.modules["skimage"], "_uwie_standin", False):
        return                                           # a real scikit-image is importable: use it
    from scipy.stats import entropy as scipy_entropy

    def shannon_entropy(image, base=2):
        _, counts = np.unique(image, return_counts=True)
        return scipy_entropy(counts, base=base)

    sk = types.ModuleType("skimage")
    sk._uwie_standin = True
    for sub in ("color", "exposure", "measure", "feature", "filters"):
        m = types.ModuleType("skimage." + sub)
        setattr(sk, sub, m)
        sys.modules["skimage." + sub] = m
    sk.measure.shannon_entropy = shannon_entropy
    sys.modules["skimage"] = sk


def install_plot_standins() -> None:
    """main.py imports matplotlib / seaborn for plotting only (main.py:12-13); absent here."""
    for name in ("matplotlib", "matplotlib.pyplot", "seaborn"):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:                            # noqa: BLE001
                sys.modules[name] = types.ModuleType(name)
    if not hasattr(sys.modules["matplotlib"], "pyplot"):
        sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]


def load_reference(which=("config", "enhancement_strategies", "quality_assessment", "six_stadigy")):
    """Import the reference's own modules (unmodified) and return them as a dict, or None if absent."""
    d = reference_dir()
    if d is None:
        return None
    try:
        import skimage  # noqa: F401
    except Exception:                                    # noqa: BLE001
        install_skimage_standin()
    if d not in sys.path:
        sys.path.insert(0, d)
    mods = {}
    for name in which:
        mods[name] = __import__(name)
        if os.path.dirname(os.path.abspath(mods[name].__file__)) != os.path.abspath(d):
            raise ImportError(f"{name} resolved to {mods[name].__file__}, not to the reference in {d}")
    mods["_dir"] = d
    return mods


def p6_frame(mods, frame: np.ndarray):
    """six_stadigy.py:406-431 loop body + comprehensive_assessment of each output, by the reference itself."""
    ss, qa = mods["six_stadigy"], mods["quality_assessment"]
    es = ss.EnhancementStrategies
    img = frame.astype(np.float32) / 255.0
    kind = ss.detect_image_type(img)
    cur = ss.color_correction(img, kind)
    outs, scores = [], []
    for fn in (es.strategy1_strong_dehazing, es.strategy2_medium_dehazing, es.strategy3_light_dehazing,
               es.strategy4_clahe_enhancement, es.strategy5_white_balance, es.strategy6_histogram_eq):
        enh = fn(cur)
        outs.append((enh * 255).astype(np.uint8))
        total, sub = qa.QualityAssessment.comprehensive_assessment(enh)
        scores.append([total] + [sub[k] for k in QA_KEYS])
    return {"type": kind, "out": np.stack(outs), "scores": np.asarray(scores, np.float64)}


def p5_frame(mods, frame: np.ndarray):
    """main.py:108-155 loop body (every Config.STRATEGIES entry, QA with Config.QUALITY_WEIGHTS, first maximum)."""
    cfg, es, qa = mods["config"], mods["enhancement_strategies"], mods["quality_assessment"]
    img = frame.astype(np.float32) / 255.0
    outs, scores = [], []
    for key, params in cfg.Config.STRATEGIES.items():
        enh = es.EnhancementStrategies.apply_strategy(img, key, params)
        outs.append((enh * 255).astype(np.uint8))
        total, sub = qa.QualityAssessment.comprehensive_assessment(enh, weights=cfg.Config.QUALITY_WEIGHTS)
        scores.append([total] + [sub[k] for k in QA_KEYS])
    scores = np.asarray(scores, np.float64)
    best = int(np.argmax(scores[:, 0]))
    return {"out": np.stack(outs), "scores": scores, "best": best}
