"""CPU oracle for the per-image enhance + score hot path.  TEST INFRASTRUCTURE ONLY.

This module is the checker the CUDA path is compared against.  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import it;
the product package never does (it fails loudly when the CUDA library is missing).

It restates, in NumPy + OpenCV (the libraries the reference itself calls), the arithmetic of
the reference's three hot-path files.  Every function cites the reference lines it follows
(paths relative to /root/reference):

    SS = six_stadigy.py            (six hard-coded strategies, float32 image buffers)
    ES = enhancement_strategies.py (five dict-parameterised strategies, float64 internals)
    QA = quality_assessment.py     (eight no-reference sub-scores + weighted total)

Parity is PINNED: tests/golden/make_golden.py imports the unmodified reference in the build
container, runs it on seeded synthetic frames and commits the outputs under tests/golden/;
tests/test_oracle_golden.py checks this restatement against those fixtures bit for bit.
The restatement is organised as data-driven recipes over rectangles and planes rather than the
reference's class-per-library layout; only the arithmetic (operation order, dtypes, constants)
is kept, because that is what parity is defined on.
"""
from __future__ import annotations

import math

import cv2
import numpy as np
from scipy.stats import entropy as _scipy_entropy

# ----------------------------------------------------------------------------------------
# shared primitives
# ----------------------------------------------------------------------------------------


def to_unit_f32(frame_u8: np.ndarray) -> np.ndarray:
    """uint8 RGB -> float32 in [0, 1]  (SS:406, main.py:108)."""
    return frame_u8.astype(np.float32) / 255.0


def quantise_u8(img: np.ndarray) -> np.ndarray:
    """Truncating float -> uint8 used before every OpenCV call (ES:298, SS:204, QA:25 ...)."""
    return (img * 255).astype(np.uint8)


def gray_u8(img: np.ndarray) -> np.ndarray:
    return cv2.cvtColor(quantise_u8(img), cv2.COLOR_RGB2GRAY)


def box_mean(plane: np.ndarray, r: int) -> np.ndarray:
    """Normalised r x r box (NOT 2r+1), anchor r//2, BORDER_REFLECT_101 (ES:31, SS:31)."""
    return cv2.boxFilter(plane, cv2.CV_64F, (r, r))


def guided_filter(guide, src, r, eps):
    """He et al. guided filter in float64 (ES:16-46, SS:25-46; identical in both)."""
    guide = guide.astype(np.float64)
    src = src.astype(np.float64)
    m_g = box_mean(guide, r)
    m_s = box_mean(src, r)
    cov = box_mean(guide * src, r) - m_g * m_s
    var = box_mean(guide * guide, r) - m_g * m_g
    a = cov / (var + eps)
    b = m_s - a * m_g
    return box_mean(a, r) * guide + box_mean(b, r)


def canny_count(gray: np.ndarray):
    """Number of edge pixels of cv2.Canny(gray, 50, 150) (ES:181, SS:150, QA:139).

    Kept as np.int64 on purpose: count / n is then a float64 *array scalar*, which promotes the
    float32 partial sums it is combined with to float64, exactly as in the reference."""
    return np.sum(cv2.Canny(gray, 50, 150) > 0)


def region_score(block: np.ndarray) -> float:
    """Quadtree score Q of one block (ES:147-188, SS:115-157).

    Q = mean brightness + (sumB + sumG - 2 sumR)/n - mean channel variance - Canny edge fraction.
    Sums run in the block's own dtype, exactly as the reference's np.sum / np.mean calls do.
    """
    rows, cols, _ = block.shape
    n = rows * cols
    cr, cg, cb = block[:, :, 0], block[:, :, 1], block[:, :, 2]
    s_r, s_g, s_b = np.sum(cr), np.sum(cg), np.sum(cb)
    brightness = (s_r + s_g + s_b) / (3 * n)
    cast = (s_b + s_g - 2 * s_r) / n
    spread = (np.sum((cr - np.mean(cr)) ** 2) / n
              + np.sum((cg - np.mean(cg)) ** 2) / n
              + np.sum((cb - np.mean(cb)) ** 2) / n) / 3
    edges = canny_count(cv2.cvtColor(quantise_u8(block), cv2.COLOR_RGB2GRAY)) / n
    return brightness + cast - spread - edges


def quadtree_airlight(img: np.ndarray, min_size: int = 1, trace: list | None = None):
    """Atmospheric light by quadtree descent (ES:77-144, SS:48-113).

    Split the current rectangle at (rows//2, cols//2), score the four quadrants, descend into the
    FIRST maximum, stop when rows <= min_size or cols <= min_size; the light is the first pixel
    with the largest R+G+B in the final rectangle (ES:191-206, SS:159-165).
    Returns the (3,) pixel (a copy).  `trace`, when given, receives one
    (y0, y1, x0, x1, [Q0..Q3]) tuple per level so tests can report decision margins.
    """
    y0, x0 = 0, 0
    y1, x1 = img.shape[0], img.shape[1]
    while (y1 - y0) > min_size and (x1 - x0) > min_size:
        ym = y0 + (y1 - y0) // 2
        xm = x0 + (x1 - x0) // 2
        rects = ((y0, ym, x0, xm), (y0, ym, xm, x1), (ym, y1, x0, xm), (ym, y1, xm, x1))
        q = [region_score(img[a:b, c:d, :]) for (a, b, c, d) in rects]
        if trace is not None:
            trace.append((y0, y1, x0, x1, [float(v) for v in q]))
        y0, y1, x0, x1 = rects[int(np.argmax(q))]
    block = img[y0:y1, x0:x1, :]
    flat = int(np.argmax(np.sum(block, axis=2)))
    ry, rx = np.unravel_index(flat, block.shape[:2])
    return block[ry, rx, :].copy()


def percentile_stretch(img: np.ndarray, lo, hi, guard: float) -> np.ndarray:
    """Per-channel clip((x - p_lo) / (p_hi - p_lo + guard), 0, 1) with np.percentile (linear).

    ES:252-273 (guard 1e-10, float64 data), SS:190-199 and SS:210-219 (guard 1e-6, float32 data;
    white_balance(p) is this function with lo = p, hi = 100 - p).  Output dtype = input dtype.
    """
    out = np.zeros_like(img)
    for c in range(3):
        ch = img[:, :, c]
        p_lo = np.percentile(ch, lo)
        p_hi = np.percentile(ch, hi)
        out[:, :, c] = np.clip((ch - p_lo) / (p_hi - p_lo + guard), 0, 1)
    return out


def clahe_lab_u8(img: np.ndarray, clip_limit: float, grid=(8, 8)) -> np.ndarray:
    """CLAHE on the 8-bit Lab L plane, returned as uint8 RGB (ES:288-306, SS:201-207)."""
    lab = cv2.cvtColor(quantise_u8(img), cv2.COLOR_RGB2LAB)
    lab[:, :, 0] = cv2.createCLAHE(clipLimit=clip_limit, tileGridSize=grid).apply(lab[:, :, 0])
    return cv2.cvtColor(lab, cv2.COLOR_LAB2RGB)


# ----------------------------------------------------------------------------------------
# SS: six_stadigy.py
# ----------------------------------------------------------------------------------------

IMAGE_TYPES = ("normal", "greenish", "bluish")


def ss_image_type(img: np.ndarray) -> str:
    """Colour-cast class from the channel means (SS:292-302)."""
    r, g, b = img.mean(axis=(0, 1))
    if g > r and g > b and (g - r) > 0.05:
        return "greenish"
    if b > r and b > g and (b - r) > 0.05:
        return "bluish"
    return "normal"


def ss_correct_cast(img: np.ndarray, kind: str) -> np.ndarray:
    """Scale the dominant channel by 0.85 and clip (SS:305-323)."""
    if kind == "normal":
        return img
    ch = 1 if kind == "greenish" else 2
    out = img.copy()
    out[:, :, ch] = out[:, :, ch] * 0.85
    return np.clip(out, 0, 1)


def ss_transmission(img, light, omega, r, eps):
    """t = clip(GF(gray, clip(1 - omega * min_c(img / (A + 1e-6)), .1, 1)), .1, 1)  (SS:167-180)."""
    dark = np.min(img / (light.reshape(1, 1, 3) + 1e-6), axis=2)
    t0 = np.clip(1 - omega * dark, 0.1, 1.0)
    guide = gray_u8(img).astype(np.float64) / 255.0
    return np.clip(guided_filter(guide, t0, r, eps), 0.1, 1.0)


def ss_restore(img, light, t):
    """(I - A) / t + A per channel, stored into a float32 buffer, clipped (SS:182-188)."""
    out = np.zeros_like(img)
    for c in range(3):
        out[:, :, c] = (img[:, :, c] - light[c]) / t + light[c]
    return np.clip(out, 0, 1)


def ss_clahe(img, clip_limit):
    return clahe_lab_u8(img, clip_limit).astype(np.float32) / 255.0


# (name, [steps]); the three dehaze recipes share one air-light estimate per frame because
# SS:233/244/254 call estimate_atmospheric_light with identical arguments.
SS_RECIPES = (
    ("strong_dehazing", (("dehaze", 0.3, 20, 5e-1), ("stretch", 5, 98), ("clahe", 3.0), ("gamma", 1.5))),
    ("medium_dehazing", (("dehaze", 0.5, 15, 5e-1), ("stretch", 15, 95), ("clahe", 2.0))),
    ("light_dehazing", (("dehaze", 0.7, 10, 1e-1), ("stretch", 20, 85), ("stretch", 2, 98))),
    ("clahe_enhancement", (("clahe", 4.0), ("stretch", 10, 95), ("stretch", 3, 97), ("gamma", 1.3))),
    ("white_balance", (("stretch", 2, 98), ("stretch", 15, 90), ("clahe", 1.5), ("gamma", 1.2))),
    ("histogram_eq", (("stretch", 5, 98), ("clahe", 3.5), ("gamma", 1.4))),
)
SS_NAMES = tuple(name for name, _ in SS_RECIPES)


def ss_run_recipe(img: np.ndarray, steps, light=None) -> np.ndarray:
    """One of strategy1..6 (SS:230-285) on a float32 frame."""
    cur = img
    for step in steps:
        op = step[0]
        if op == "dehaze":
            _, omega, r, eps = step
            if light is None:
                light = quadtree_airlight(cur, 1)
            t = ss_transmission(cur, light, omega, r, eps)
            cur = ss_restore(cur, light, t)
        elif op == "stretch":
            cur = percentile_stretch(cur, step[1], step[2], 1e-6)
        elif op == "clahe":
            cur = ss_clahe(cur, step[1])
        elif op == "gamma":
            cur = np.power(cur, step[1])  # SS:221-224: img ** gamma, no clip
        else:  # pragma: no cover
            raise KeyError(op)
    return cur


def ss_strategy(img: np.ndarray, index: int) -> np.ndarray:
    return ss_run_recipe(img, SS_RECIPES[index][1])


# ----------------------------------------------------------------------------------------
# ES: enhancement_strategies.py
# ----------------------------------------------------------------------------------------

ES_DEFAULT_STRATEGIES = {  # config.py:28-75, kept as data
    "strong_dehazing": {"name": "StrongDehazing", "omega": 0.5, "guided_radius": 15, "L_low": 10,
                        "L_high": 95, "gamma": 1.2, "apply_gamma": True},
    "medium_dehazing": {"name": "MediumDehazing", "omega": 0.6, "guided_radius": 20, "L_low": 15,
                        "L_high": 92, "apply_gamma": True},
    "light_enhancement": {"name": "LightEnhancement", "omega": 0.4, "guided_radius": 10, "L_low": 15,
                          "L_high": 95, "apply_gamma": False},
    "clahe_enhancement": {"name": "CLAHEEnhancement", "clip_limit": 2.0, "tile_grid_size": (8, 8),
                          "apply_gamma": False},
    "histogram_equalization": {"name": "HistogramEqualization", "L_low": 10, "L_high": 95},
}
ES_QUALITY_WEIGHTS = {"contrast": 0.25, "sharpness": 0.20, "entropy": 0.15, "saturation": 0.15,
                      "brightness": 0.15, "edge_density": 0.10}  # config.py:78-85

# per-strategy fall-back parameters when a key is missing from the dict (ES:356-357, 364-365 ...)
_ES_DEFAULTS = {
    "strong_dehazing": dict(omega=0.5, guided_radius=15, L_low=10, L_high=95),
    "medium_dehazing": dict(omega=0.6, guided_radius=20, L_low=15, L_high=92),
    "light_enhancement": dict(omega=0.4, guided_radius=10, L_low=15, L_high=95),
    "clahe_enhancement": dict(clip_limit=2.0, tile_grid_size=(8, 8), L_low=20, L_high=85),
    "histogram_equalization": dict(L_low=10, L_high=95),
}


def es_transmission(img, light_plane, omega=0.95, r=15, eps=1e-3):
    """ES:209-234: t0 is NOT clipped before the guided filter; the guard is 1e-10."""
    dark = np.min(img / (light_plane + 1e-10), axis=2)
    t0 = 1 - omega * dark
    guide = gray_u8(img).astype(np.float64) / 255.0
    return np.clip(guided_filter(guide, t0, r, eps), 0.1, 1.0)


def es_recover(img, t, light_plane):
    """ES:237-249."""
    return np.clip((img - light_plane) / np.expand_dims(t, axis=2) + light_plane, 0, 1)


def es_gamma(img, gamma=1.2):
    """ES:276-285: x ** (1 / gamma), clipped."""
    return np.clip(np.power(img, 1.0 / gamma), 0, 1)


def es_equalize(img):
    """Per-channel cv2.equalizeHist on the truncated uint8 frame (ES:331-345)."""
    q = quantise_u8(img)
    out = np.zeros_like(q)
    for c in range(3):
        out[:, :, c] = cv2.equalizeHist(q[:, :, c])
    return out.astype(np.float64) / 255.0


def es_strategy(img: np.ndarray, name: str, params: dict, light=None) -> np.ndarray:
    """apply_strong/medium/light/clahe/histogram_* (ES:349-474).  Raises KeyError on unknown names."""
    d = _ES_DEFAULTS[name]

    def get(key):
        return params.get(key, d[key])

    if name in ("strong_dehazing", "medium_dehazing", "light_enhancement"):
        if light is None:
            light = quadtree_airlight(img, 1)
        plane = np.tile(light.reshape(1, 1, 3), (img.shape[0], img.shape[1], 1))  # ES:142
        t = es_transmission(img, plane, omega=get("omega"), r=get("guided_radius"))
        cur = es_recover(img, t, plane)
    elif name == "clahe_enhancement":
        cur = clahe_lab_u8(img, get("clip_limit"), get("tile_grid_size")).astype(np.float64) / 255.0
    elif name == "histogram_equalization":
        cur = es_equalize(img)
    else:
        raise KeyError(name)
    cur = percentile_stretch(cur, get("L_low"), get("L_high"), 1e-10)
    if params.get("apply_gamma", False):
        cur = es_gamma(cur, params.get("gamma", 1.2))
    return cur


# ----------------------------------------------------------------------------------------
# QA: quality_assessment.py
# ----------------------------------------------------------------------------------------

QA_KEYS = ("contrast", "sharpness", "entropy", "saturation", "brightness", "edge_density",
           "colorfulness", "naturalness")
QA_DEFAULT_WEIGHTS = {"contrast": 0.20, "sharpness": 0.20, "entropy": 0.15, "saturation": 0.15,
                      "brightness": 0.10, "edge_density": 0.10, "colorfulness": 0.05,
                      "naturalness": 0.05}  # QA:229-238


def _value_entropy(plane: np.ndarray) -> float:
    """skimage.measure.shannon_entropy (>= 0.19): scipy entropy of the value counts, base 2."""
    _, counts = np.unique(plane, return_counts=True)
    return float(_scipy_entropy(counts, base=2))


def qa_sharpness(img) -> float:
    """QA:36-55.  With OpenCV 4.x cv2.Laplacian(float32, CV_64F) raises, and QA:250-251 turns the
    failure into the constant 50.0 — so this returns 50.0 whenever the call fails."""
    try:
        g = gray_u8(img).astype(np.float32) / 255.0
        lap = cv2.Laplacian(g, cv2.CV_64F)
        return float(np.clip(np.var(lap) / 0.5 * 100, 0, 100))
    except Exception:
        return 50.0


def qa_scores(img: np.ndarray) -> dict:
    """The eight sub-scores of QA:14-212 for a float RGB frame in [0, 1]."""
    q = quantise_u8(img)
    g8 = cv2.cvtColor(q, cv2.COLOR_RGB2GRAY)
    g = g8.astype(np.float32) / 255.0
    hsv = cv2.cvtColor(q, cv2.COLOR_RGB2HSV).astype(np.float32) / 255.0
    sat = hsv[:, :, 1]
    lum = cv2.cvtColor(q, cv2.COLOR_RGB2LAB).astype(np.float32)[:, :, 0]
    out = {}
    out["contrast"] = np.clip(np.std(g) / 0.5 * 100, 0, 100)                       # QA:25-31
    out["sharpness"] = qa_sharpness(img)                                           # QA:46-53
    out["entropy"] = np.clip((_value_entropy(g) - 4) / 4 * 100, 0, 100)            # QA:68-73
    out["saturation"] = np.clip(np.mean(sat) * 100, 0, 100)                        # QA:88-95
    out["brightness"] = 100 - np.clip(abs(np.mean(lum) - 128) / 128 * 100, 0, 100)  # QA:111-121
    out["edge_density"] = np.clip(canny_count(g8) / g8.size / 0.2 * 100, 0, 100)   # QA:136-145
    r, gg, b = img[:, :, 0], img[:, :, 1], img[:, :, 2]                            # QA:161-182
    rg = r - gg
    yb = 0.5 * (r + gg) - b
    spread = np.sqrt(np.std(rg) ** 2 + np.std(yb) ** 2)
    offset = np.sqrt(np.mean(rg) ** 2 + np.mean(yb) ** 2)
    out["colorfulness"] = np.clip((spread + 0.3 * offset) / 0.5 * 100, 0, 100)
    bad = (np.sum(sat > 0.9) / sat.size + np.sum(g < 0.1) / g.size + np.sum(g > 0.9) / g.size)
    out["naturalness"] = 100 - np.clip(bad * 200, 0, 100)                          # QA:198-210
    return out


def qa_total(img: np.ndarray, weights: dict | None = None):
    """comprehensive_assessment (QA:215-286): (weighted total, sub-score dict)."""
    w = QA_DEFAULT_WEIGHTS if weights is None else weights
    s = qa_scores(img)
    return sum(s[k] * w.get(k, 0) for k in s), s


def scores_row(total, sub) -> np.ndarray:
    """[total, 8 sub-scores in QA_KEYS order] as float64 — the record layout of the CUDA path."""
    return np.asarray([total] + [sub[k] for k in QA_KEYS], dtype=np.float64)


# ----------------------------------------------------------------------------------------
# whole-frame pipelines (the two bench workloads of SURVEY.md §8d)
# ----------------------------------------------------------------------------------------


def p6_frame(frame_u8: np.ndarray, weights: dict | None = None, with_scores: bool = True):
    """SS per-image loop (SS:406-431) + comprehensive_assessment on each strategy output.

    Returns dict(type=int, light=(3,), out=uint8[6,H,W,3], scores=f64[6,9], best=int)."""
    img = to_unit_f32(frame_u8)
    kind = ss_image_type(img)
    cur = ss_correct_cast(img, kind)
    light = quadtree_airlight(cur, 1)
    outs, rows = [], []
    for _, steps in SS_RECIPES:
        enh = ss_run_recipe(cur, steps, light=light)
        outs.append(quantise_u8(enh))                        # SS:430
        if with_scores:
            rows.append(scores_row(*qa_total(enh, weights)))
    res = {"type": IMAGE_TYPES.index(kind), "light": np.asarray(light, dtype=np.float32),
           "out": np.stack(outs)}
    if with_scores:
        res["scores"] = np.stack(rows)
        res["best"] = int(np.argmax(res["scores"][:, 0]))
    return res


def p5_frame(frame_u8: np.ndarray, strategies: dict | None = None, weights: dict | None = None):
    """main.py:108-155 loop body (without FeatureExtractor): five ES strategies, QA each, first max.

    Returns dict(out=uint8[S,H,W,3], scores=f64[S,9], best=int, light=(3,))."""
    strategies = ES_DEFAULT_STRATEGIES if strategies is None else strategies
    weights = ES_QUALITY_WEIGHTS if weights is None else weights
    img = to_unit_f32(frame_u8)
    light = quadtree_airlight(img, 1)
    outs, rows = [], []
    for key, params in strategies.items():
        enh = es_strategy(img, key, params, light=light)
        outs.append(quantise_u8(enh))                        # main.py:155
        rows.append(scores_row(*qa_total(enh, weights)))
    scores = np.stack(rows)
    return {"out": np.stack(outs), "scores": scores, "best": int(np.argmax(scores[:, 0])),
            "light": np.asarray(light, dtype=np.float32)}


def np_percentile_plan(n: int, q, f32: bool):
    """(k, gamma) NumPy's linear method uses for percentile q of n values.

    For float32 data NumPy 2.x divides q by float32(100) and evaluates the virtual index
    (n - 1) * q in float32 (numpy/lib/_function_base_impl.py, percentile and
    _QuantileMethods['linear']); for float64 data everything is float64.  Exposed so host-logic tests
    can pin the C++ planner against it."""
    if f32:
        qq = np.true_divide(q, np.float32(100))
    else:
        qq = np.true_divide(q, 100)
    vi = (n - 1) * qq
    k = math.floor(float(vi))
    if float(vi) >= n - 1:
        return n - 1, type(vi)(0)
    return int(k), vi - type(vi)(k)


# ----------------------------------------------------------------------------------------
# VGG-guided path: the per-pixel pieces around the network
#   VGG = vgg_16_UIE.py, UTM = use_trained_model.py
# ----------------------------------------------------------------------------------------


def vgg_stretch_rank(level: float, n: int) -> int:
    """int((L / 100.0) * n) clamped to [0, n - 1]  (VGG:78-82; L is a float32 tensor item)."""
    k = int((float(np.float32(level)) / 100.0) * n)
    return max(0, min(k, n - 1))


def vgg_enhance(img: np.ndarray, l_low, l_high, omega=None, gamma=None) -> np.ndarray:
    """DifferentiableEnhancement.forward on a float32 NCHW batch (VGG:32-55), all steps in float32."""
    img = np.asarray(img, dtype=np.float32)
    B, C, H, W = img.shape
    n = H * W
    out = np.zeros_like(img)
    for b in range(B):
        for c in range(C):                                           # VGG:70-90
            plane = img[b, c]
            srt = np.sort(plane.reshape(-1))
            p_lo = srt[vgg_stretch_rank(l_low[b], n)]
            p_hi = srt[vgg_stretch_rank(l_high[b], n)]
            rng = np.float32(np.float32(p_hi - p_lo) + np.float32(1e-8))
            out[b, c] = np.clip((plane - p_lo) / rng, 0, 1)
    if omega is not None:                                            # VGG:94-117
        w = np.asarray(omega, dtype=np.float32).reshape(-1, 1, 1, 1)
        dark = out.min(axis=1, keepdims=True)
        t = np.clip(np.float32(1) - w * dark, np.float32(0.1), np.float32(1.0))
        out = np.clip((out - np.float32(0.6)) / t + np.float32(0.6), 0, 1).astype(np.float32)
    if gamma is not None:                                            # VGG:119-128
        g = np.asarray(gamma, dtype=np.float32).reshape(-1, 1, 1, 1)
        out = np.power(out + np.float32(1e-8), g).astype(np.float32)
    return np.clip(out, 0, 1).astype(np.float32)                     # VGG:55


def vgg_features(img: np.ndarray) -> np.ndarray:
    """extract_all_features (VGG:435-466): 18 statistics, zero-padded to 79, float32."""
    img = np.asarray(img)
    if img.dtype == np.uint8:                                        # VGG:425-433
        img = img.astype(np.float32) / 255.0
    else:
        img = img.astype(np.float32)
        if img.max() > 1.0:
            img = img / 255.0
    feats = []
    for c in range(3):
        ch = img[:, :, c]
        feats += [float(np.mean(ch)), float(np.std(ch)), float(np.min(ch)), float(np.max(ch)),
                  float(np.median(ch))]
    feats += [float(np.mean(img)), float(np.std(img)), float(np.mean(img ** 2))]
    feats += [0.0] * (79 - len(feats))
    return np.array(feats[:79], dtype=np.float32)


def vgg_preprocess(img: np.ndarray, size: int = 224) -> np.ndarray:
    """_preprocess_for_vgg (UTM:39-46): 8-bit bilinear resize, /255, ImageNet normalisation, CHW."""
    if img.dtype != np.uint8:
        img = (img * 255).astype(np.uint8)
    small = cv2.resize(img, (size, size), interpolation=cv2.INTER_LINEAR).astype(np.float32) / 255.0
    chw = np.ascontiguousarray(small.transpose(2, 0, 1))
    mean = np.array([0.485, 0.456, 0.406], np.float32).reshape(3, 1, 1)
    std = np.array([0.229, 0.224, 0.225], np.float32).reshape(3, 1, 1)
    return ((chw - mean) / std).astype(np.float32)


# ----------------------------------------------------------------------------------------
# FE = feature_extraction.py: the 79 hand-crafted features
# Order: colour 35 | texture 22 | frequency 5 | edge 7 | quality 10   (FE:249-297)
# The three scikit-image functions of the texture group are restated from their documented
# algorithm; scikit-image is not installable here, so that group is UNPINNED.
# ----------------------------------------------------------------------------------------


def sk_local_binary_pattern_uniform(image: np.ndarray, P: int = 8, R: float = 1.0) -> np.ndarray:
    """skimage.feature.local_binary_pattern(image, P, R, method='uniform') (FE:97), float64 output.

    Taps at (-R sin(2 pi p / P), R cos(2 pi p / P)) rounded to 5 decimals; bilinear interpolation in
    double with constant 0 outside the image; bit = tap >= centre; code = number of set bits when the
    bit string has at most two transitions between its P - 1 adjacent pairs (not circular), else P + 1."""
    img = np.ascontiguousarray(image, dtype=np.float64)
    rows, cols = img.shape
    ang = 2.0 * np.pi * np.arange(P, dtype=np.float64) / P
    rp = np.round(-R * np.sin(ang), 5)
    cp = np.round(R * np.cos(ang), 5)
    pad = np.zeros((rows + 4, cols + 4), np.float64)
    pad[2:-2, 2:-2] = img
    rr, cc = np.meshgrid(np.arange(rows, dtype=np.float64), np.arange(cols, dtype=np.float64), indexing="ij")
    bits = np.zeros((P, rows, cols), np.int64)
    for p in range(P):
        r, c = rr + rp[p], cc + cp[p]
        r0, c0 = np.floor(r), np.floor(c)
        dr, dc = r - r0, c - c0
        i0, j0 = r0.astype(np.int64) + 2, c0.astype(np.int64) + 2
        tl, tr = pad[i0, j0], pad[i0, j0 + 1]
        bl, br = pad[i0 + 1, j0], pad[i0 + 1, j0 + 1]
        top = (1 - dc) * tl + dc * tr
        bot = (1 - dc) * bl + dc * br
        bits[p] = ((1 - dr) * top + dr * bot - img) >= 0
    changes = np.zeros((rows, cols), np.int64)
    for p in range(P - 1):
        changes += bits[p] != bits[p + 1]
    return np.where(changes <= 2, bits.sum(axis=0), P + 1).astype(np.float64)


_GLCM_OFFSETS = ((0, 1), (1, 1), (1, 0), (1, -1))   # (round(sin a), round(cos a)) for a = 0, pi/4, pi/2, 3pi/4


def sk_graycomatrix(image: np.ndarray, levels: int = 256) -> np.ndarray:
    """graycomatrix(image, [1], [0, pi/4, pi/2, 3pi/4], levels, symmetric=True, normed=True) (FE:107-112):
    float64 [levels, levels, 1, 4]; offset of angle a is (round(sin a), round(cos a)) with rows counted
    downwards, in-bounds pairs only."""
    img = np.asarray(image)
    rows, cols = img.shape
    out = np.zeros((levels, levels, 1, 4), np.float64)
    for k, (dr, dc) in enumerate(_GLCM_OFFSETS):
        r0, r1 = max(0, -dr), rows - max(0, dr)
        c0, c1 = max(0, -dc), cols - max(0, dc)
        a = img[r0:r1, c0:c1].reshape(-1).astype(np.int64)
        b = img[r0 + dr:r1 + dr, c0 + dc:c1 + dc].reshape(-1).astype(np.int64)
        m = np.bincount(a * levels + b, minlength=levels * levels).reshape(levels, levels).astype(np.float64)
        m = m + m.T
        out[:, :, 0, k] = m / m.sum()
    return out


def sk_graycoprops(P: np.ndarray, prop: str) -> np.ndarray:
    """graycoprops (FE:117): [1, 4] array of one property per (distance, angle)."""
    levels = P.shape[0]
    I, J = np.ogrid[0:levels, 0:levels]
    I = I.reshape(levels, 1, 1, 1).astype(np.float64)
    J = J.reshape(1, levels, 1, 1).astype(np.float64)
    if prop == "contrast":
        return np.sum(P * (I - J) ** 2, axis=(0, 1))
    if prop == "dissimilarity":
        return np.sum(P * np.abs(I - J), axis=(0, 1))
    if prop == "homogeneity":
        return np.sum(P / (1.0 + (I - J) ** 2), axis=(0, 1))
    asm = np.sum(P ** 2, axis=(0, 1))
    if prop == "ASM":
        return asm
    if prop == "energy":
        return np.sqrt(asm)
    if prop == "correlation":
        mi = np.sum(I * P, axis=(0, 1))
        mj = np.sum(J * P, axis=(0, 1))
        si = np.sqrt(np.sum(P * (I - mi) ** 2, axis=(0, 1)))
        sj = np.sqrt(np.sum(P * (J - mj) ** 2, axis=(0, 1)))
        cov = np.sum(P * (I - mi) * (J - mj), axis=(0, 1))
        res = np.ones_like(cov)
        ok = ~((si < 1e-15) | (sj < 1e-15))
        res[ok] = cov[ok] / (si[ok] * sj[ok])
        return res
    raise ValueError(prop)


GLCM_PROPS = ("contrast", "dissimilarity", "homogeneity", "energy", "correlation", "ASM")   # FE:115


def _moments4(x: np.ndarray):
    """mean, std, skew, kurtosis the way FE:33-40 asks for them (np / scipy.stats on a float32 vector)."""
    from scipy import stats
    return [np.mean(x), np.std(x), stats.skew(x), stats.kurtosis(x)]


def fe_color(img: np.ndarray) -> np.ndarray:
    """extract_color_features (FE:17-77): 35 values."""
    q = quantise_u8(img)
    lab = cv2.cvtColor(q, cv2.COLOR_RGB2LAB).astype(np.float32)
    hsv = cv2.cvtColor(q, cv2.COLOR_RGB2HSV).astype(np.float32)
    out = []
    for c in range(3):
        out += _moments4(lab[:, :, c].reshape(-1))
    for c in range(3):
        ch = hsv[:, :, c].reshape(-1)
        out += [np.mean(ch), np.std(ch)]
    a, b = lab[:, :, 1], lab[:, :, 2]
    ma, mb = np.mean(a), np.mean(b)
    M = np.sqrt(ma ** 2 + mb ** 2)
    D = np.sqrt(np.mean(np.abs(a - ma)) ** 2 + np.mean(np.abs(b - mb)) ** 2)
    out += [M / (D + 1e-10), M, D, ma, mb]
    for c in range(3):
        ch = img[:, :, c].reshape(-1)
        out += [np.mean(ch), np.std(ch), np.min(ch), np.max(ch)]
    return np.array(out)


def fe_texture(img: np.ndarray) -> np.ndarray:
    """extract_texture_features (FE:80-120): LBP histogram (10) + GLCM statistics (12)."""
    gray = gray_u8(img).astype(np.float32) / 255.0
    g8 = (gray * 255).astype(np.uint8)
    lbp = sk_local_binary_pattern_uniform(g8, 8, 1)
    hist, _ = np.histogram(lbp.ravel(), bins=10, range=(0, 10), density=True)
    small = cv2.resize(g8, (128, 128))
    P = sk_graycomatrix(small, 256)
    out = list(hist)
    for prop in GLCM_PROPS:
        v = sk_graycoprops(P, prop).reshape(-1)
        out += [np.mean(v), np.std(v)]
    return np.array(out)


def fe_frequency(img: np.ndarray) -> np.ndarray:
    """extract_frequency_features (FE:123-158): DCT energy split + |DCT| statistics."""
    d = cv2.dct(gray_u8(img).astype(np.float32))
    h, w = d.shape
    total = np.sum(d ** 2)
    return np.array([np.sum(d[:h // 4, :w // 4] ** 2) / total,
                     np.sum(d[h // 4:h // 2, w // 4:w // 2] ** 2) / total,
                     np.sum(d[h // 2:, w // 2:] ** 2) / total,
                     np.mean(np.abs(d)), np.std(np.abs(d))])


def fe_edge(img: np.ndarray) -> np.ndarray:
    """extract_edge_features (FE:161-200)."""
    gray = gray_u8(img).astype(np.float32) / 255.0
    sx = cv2.Sobel(gray, cv2.CV_32F, 1, 0, ksize=3)
    sy = cv2.Sobel(gray, cv2.CV_32F, 0, 1, ksize=3)
    mag = np.sqrt(sx ** 2 + sy ** 2)
    g8 = (gray * 255).astype(np.uint8)
    edges = cv2.Canny(g8, 50, 150)
    lap = cv2.Laplacian(g8, cv2.CV_64F, ksize=3)
    return np.array([np.mean(mag), np.std(mag), np.max(mag), np.sum(edges > 0) / edges.size,
                     np.mean(np.abs(lap)), np.std(lap), np.var(lap)])


def fe_quality(img: np.ndarray) -> np.ndarray:
    """extract_quality_features (FE:203-246)."""
    q = quantise_u8(img)
    gray = cv2.cvtColor(q, cv2.COLOR_RGB2GRAY).astype(np.float32) / 255.0
    sat = cv2.cvtColor(q, cv2.COLOR_RGB2HSV).astype(np.float32)[:, :, 1] / 255.0
    return np.array([np.std(gray), _value_entropy(gray), np.mean(gray), np.median(gray),
                     np.percentile(gray, 25), np.percentile(gray, 75), np.max(gray) - np.min(gray),
                     np.mean(sat), np.std(sat), np.sqrt(np.mean((gray - np.mean(gray)) ** 2))])


FE_GROUPS = (("color", fe_color, 35), ("texture", fe_texture, 22), ("frequency", fe_frequency, 5),
             ("edge", fe_edge, 7), ("quality", fe_quality, 10))


def fe_all(img: np.ndarray) -> np.ndarray:
    """extract_all_features (FE:249-297): concatenation of the five groups, float64 [79]."""
    if img.dtype == np.uint8:
        img = to_unit_f32(img)
    return np.concatenate([np.asarray(fn(img), dtype=np.float64) for _, fn, _ in FE_GROUPS])
