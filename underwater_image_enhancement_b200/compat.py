"""Drop-in faces of the reference's entry points for the hot path, with the reference's names, argument
meaning and error behaviour.  INTEGRATION.md shows how the reference's own modules
bind to them.

    es_apply_strategy(img, name, params)            enhancement_strategies.py:477-508
    enhance_and_select(img, strategies, weights)    main.py:116-155 (every strategy, score, first maximum)
    detect_image_type(img)                          six_stadigy.py:292-302
    six_strategies(img)                             six_stadigy.py:409-431 (colour cast + strategy1..6)
    process_all_images_all_strategies(in, out)      six_stadigy.py:326-470 (file loop, same output names)

Images are the reference's: RGB, float in [0, 1] made from 8-bit files (imread / 255, main.py:108,
six_stadigy.py:406) or the uint8 frames themselves.  Enhanced images come back as bytes / 255 in the
library's float type (float64 for ES, float32 for SS): identical to the reference's
(enhanced * 255).astype(uint8) outputs, and within 1/255 of its float images (north_star tolerance).
All arithmetic runs in libuwie_b200.so; there is no CPU fallback.
"""
from __future__ import annotations

from pathlib import Path

import numpy as np

from . import _lib
from .engine import ES_NAMES, IMAGE_TYPES, QA_KEYS, SS_NAMES, Engine

_ENGINES: dict = {}

# config.py:28-75 defaults of each apply_* (enhancement_strategies.py:350-474)
_ES_DEFAULTS = {
    "strong_dehazing": dict(kind="dehaze", omega=0.5, guided_radius=15, L_low=10, L_high=95),
    "medium_dehazing": dict(kind="dehaze", omega=0.6, guided_radius=20, L_low=15, L_high=92),
    "light_enhancement": dict(kind="dehaze", omega=0.4, guided_radius=10, L_low=15, L_high=95),
    "clahe_enhancement": dict(kind="clahe", clip_limit=2.0, tile_grid_size=(8, 8), L_low=20, L_high=85),
    "histogram_equalization": dict(kind="equalize", L_low=10, L_high=95),
}


def _to_frame(img) -> np.ndarray:
    """The reference's float image (or the uint8 frame) -> uint8 HWC, by the reference's own truncation."""
    a = np.asarray(img)
    if a.ndim != 3 or a.shape[2] != 3:
        raise ValueError(f"expected an (H, W, 3) RGB image, got {a.shape}")
    if a.dtype == np.uint8:
        return np.ascontiguousarray(a)
    q = (a * 255).astype(np.uint8)                         # ES:298, SS:204, QA:25 ...
    if not np.array_equal(q.astype(a.dtype) / a.dtype.type(255), a):
        raise ValueError("the hot path takes images made from 8-bit data (imread / 255): this one is not byte-valued")
    return np.ascontiguousarray(q)


def _es_strategy(name: str, params: dict) -> _lib.Strategy:
    d = _ES_DEFAULTS[name]
    g = lambda k: params.get(k, d[k])                      # noqa: E731
    s = _lib.Strategy()
    if d["kind"] == "dehaze":
        s.dehaze, s.omega, s.radius, s.eps = 1, float(g("omega")), int(g("guided_radius")), 1e-3   # ES:209 default eps
        s.n_pre = 1; s.pre_lo[0] = float(g("L_low")); s.pre_hi[0] = float(g("L_high"))
    else:
        if d["kind"] == "clahe":
            s.clahe, s.clip_limit = 1, float(g("clip_limit"))
            s.tiles_x, s.tiles_y = (int(v) for v in g("tile_grid_size"))
        else:
            s.equalize = 1
        s.n_post = 1; s.post_lo[0] = float(g("L_low")); s.post_hi[0] = float(g("L_high"))
    if params.get("apply_gamma", False):
        s.gamma_on, s.gamma = 1, float(params.get("gamma", 1.2))
    return s


def _strategy_key(s: _lib.Strategy):
    return (s.dehaze, s.omega, s.radius, s.eps, s.n_pre, tuple(s.pre_lo), tuple(s.pre_hi), s.equalize, s.clahe,
            s.clip_limit, s.tiles_x, s.tiles_y, s.n_post, tuple(s.post_lo), tuple(s.post_hi), s.gamma_on, s.gamma)


def _engine(library: str, h: int, w: int, strategies=None, weights=None) -> Engine:
    key = (library, h, w, None if strategies is None else tuple(_strategy_key(s) for s in strategies),
           None if weights is None else tuple(sorted(weights.items())))
    eng = _ENGINES.get(key)
    if eng is None:
        if len(_ENGINES) >= 8:                              # a handful of shapes at most stays resident
            _ENGINES.pop(next(iter(_ENGINES))).close()
        eng = _ENGINES[key] = Engine(library, h, w, max_batch=1, strategies=strategies, weights=weights)
    return eng


# ------------------------------------------------------------------ enhancement_strategies.py / main.py
def es_apply_strategy(img, strategy_name: str, params: dict):
    """EnhancementStrategies.apply_strategy: ValueError for an unknown name (ES:500-501); any failure inside
    the strategy prints and returns the input image (ES:506-508)."""
    if strategy_name not in _ES_DEFAULTS:
        raise ValueError(f"未知策略: {strategy_name}")
    try:
        frame = _to_frame(img)
        eng = _engine("es", frame.shape[0], frame.shape[1], strategies=[_es_strategy(strategy_name, params)])
        out = eng.run_host(frame[None], want_all=True)["out"][0, 0]
        return out.astype(np.float64) / 255.0
    except Exception as e:                                  # noqa: BLE001  (the reference's behaviour)
        print(f"策略 {strategy_name} 執行失敗: {e}")
        return img


def enhance_and_select(img, strategies: dict | None = None, weights: dict | None = None):
    """main.py:116-155: every strategy of `strategies` (default Config.STRATEGIES), comprehensive_assessment of
    each with `weights` (default Config.QUALITY_WEIGHTS), first maximum.
    Returns (best_name, best_image float64, {name: total}, {name: {sub-scores}})."""
    frame = _to_frame(img)
    if strategies is None:
        names, strat = list(ES_NAMES), None
    else:
        names = list(strategies)
        for n in names:
            if n not in _ES_DEFAULTS:
                raise ValueError(f"未知策略: {n}")
        strat = [_es_strategy(n, strategies[n]) for n in names]
    eng = _engine("es", frame.shape[0], frame.shape[1], strategies=strat, weights=weights)
    res = eng.run_host(frame[None], want_all=False, want_best=True)
    scores = res["scores"][0]
    best = int(res["best"][0])
    totals = {n: float(scores[i, 0]) for i, n in enumerate(names)}
    subs = {n: {k: float(scores[i, 1 + j]) for j, k in enumerate(QA_KEYS)} for i, n in enumerate(names)}
    return names[best], res["best_out"][0].astype(np.float64) / 255.0, totals, subs


# ------------------------------------------------------------------ six_stadigy.py
def six_strategies(img, with_scores: bool = False):
    """six_stadigy.py:409-431 for one image: (image_type, {strategy name: float32 image}) and, on request, the
    comprehensive_assessment rows of the six outputs."""
    frame = _to_frame(img)
    eng = _engine("six", frame.shape[0], frame.shape[1])
    res = eng.run_host(frame[None], want_all=True)
    kind = IMAGE_TYPES[int(res["type"][0])]
    outs = {n: res["out"][0, i].astype(np.float32) / np.float32(255.0) for i, n in enumerate(SS_NAMES)}
    if with_scores:
        return kind, outs, {n: res["scores"][0, i].copy() for i, n in enumerate(SS_NAMES)}
    return kind, outs


def detect_image_type(img) -> str:
    """six_stadigy.py:292-302."""
    frame = _to_frame(img)
    eng = _engine("six", frame.shape[0], frame.shape[1])
    return IMAGE_TYPES[int(eng.run_host(frame[None], want_all=False)["type"][0])]


def process_all_images_all_strategies(input_folder, output_base_folder, batch: int = 8):
    """six_stadigy.py:326-470: every image of `input_folder` through colour cast + the six strategies, written as
    <stem>_<strategy name>.png (six_stadigy.py:344-351, 430) under `output_base_folder`.  Files of one size are batched through one engine.
    Returns the log rows (filename, image_type, strategy, status, output_path) the reference writes to its CSV."""
    import cv2
    src, dst = Path(input_folder), Path(output_base_folder)
    dst.mkdir(parents=True, exist_ok=True)
    files = sorted(p for p in src.iterdir() if p.suffix.lower() in (".jpg", ".jpeg", ".png", ".tif", ".tiff"))   # SS:361
    log, groups = [], {}
    for p in files:
        bgr = cv2.imread(str(p))
        if bgr is None:
            log.append(dict(filename=p.name, image_type="", strategy="", status="failed: unreadable", output_path=""))
            continue
        groups.setdefault(bgr.shape[:2], []).append((p, cv2.cvtColor(bgr, cv2.COLOR_BGR2RGB)))
    for (h, w), items in groups.items():
        eng = Engine("six", h, w, max_batch=min(batch, len(items)))
        try:
            res = eng.run_host(np.stack([f for _, f in items]), want_all=True)
        finally:
            eng.close()
        for i, (p, _) in enumerate(items):
            kind = IMAGE_TYPES[int(res["type"][i])]
            for s, name in enumerate(SS_NAMES):
                out = dst / f"{p.stem}_{name}.png"
                ok = cv2.imwrite(str(out), cv2.cvtColor(res["out"][i, s], cv2.COLOR_RGB2BGR))
                log.append(dict(filename=p.name, image_type=kind, strategy=name,
                                status="success" if ok else "failed: write", output_path=str(out)))
    return log


def close_engines():
    while _ENGINES:
        _ENGINES.popitem()[1].close()
