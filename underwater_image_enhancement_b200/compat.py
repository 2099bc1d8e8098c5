"""The reference's entry points for the hot path, with the reference's names, argument meaning, return types and
error behaviour.  The shim modules in `shims/` (enhancement_strategies.py, six_stadigy.py,
quality_assessment.py, feature_extraction.py, use_trained_model.py) are thin re-exports of what is defined here,
so that main.py / example_usage.py / use_trained_model.py run unchanged with `shims/` first on sys.path.

Two routes, chosen per call, both on the device (there is no CPU fallback):

* byte-valued images (made from 8-bit files: imread / 255, main.py:108, six_stadigy.py:406 — every image the
  reference's own loops produce) go through the fused engine: results are bytes / 255 in the library's float
  type, identical to the reference's `(enhanced * 255).astype(uint8)` and within 1/255 of its float image;
* any other float image (example_usage.py:27 feeds np.random.rand) goes through the per-primitive operators
  (prims.py), composed exactly like the reference's apply_* / strategyN bodies.
"""
from __future__ import annotations

import time
from pathlib import Path

import numpy as np
import torch

from . import _lib, prims
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
SS_DESCRIPTIONS = ("強力去霧", "中度去霧", "輕度去霧", "CLAHE增強", "白平衡主導", "直方圖均衡")     # six_stadigy.py:344-351


# ------------------------------------------------------------------ image classification: bytes or floats
def _check_image(img) -> np.ndarray:
    a = np.asarray(img)
    if a.ndim != 3 or a.shape[2] != 3:
        raise ValueError(f"expected an (H, W, 3) RGB image, got {a.shape}")
    return a


def as_frame(img):
    """uint8 HWC frame when the image is made from bytes (k / 255 exactly, in float32 or float64, or uint8 itself),
    else None."""
    a = _check_image(img)
    if a.dtype == np.uint8:
        return np.ascontiguousarray(a)
    if a.dtype not in (np.float32, np.float64):
        a = a.astype(np.float64)
    if a.size == 0 or not (a.min() >= 0 and a.max() <= 1):
        return None
    q = (a * 255).astype(np.uint8)                         # ES:298, SS:204, QA:25 ...
    if np.array_equal(q.astype(a.dtype) / a.dtype.type(255), a):
        return np.ascontiguousarray(q)
    return None


def _to_frame(img) -> np.ndarray:
    f = as_frame(img)
    if f is None:
        raise ValueError("this entry point takes images made from 8-bit data (imread / 255)")
    return f


def _float_tensor(img) -> torch.Tensor:
    a = _check_image(img)
    if a.dtype == np.uint8:
        a = a.astype(np.float32) / np.float32(255.0)
    elif a.dtype not in (np.float32, np.float64):
        a = a.astype(np.float64)
    return prims.as_device(a)


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


def _engine(library: str, h: int, w: int, strategies=None, weights=None, colour_cast=None, max_batch=1) -> Engine:
    key = (library, h, w, None if strategies is None else tuple(_strategy_key(s) for s in strategies),
           None if weights is None else tuple(sorted(weights.items())), colour_cast, max_batch)
    eng = _ENGINES.get(key)
    if eng is None:
        if len(_ENGINES) >= 8:                              # a handful of shapes at most stays resident
            _ENGINES.pop(next(iter(_ENGINES))).close()
        lib_struct = None
        if colour_cast is not None:
            lib_struct, _ = _lib.preset(_lib.LIB_SIX if library == "six" else _lib.LIB_ES)
            lib_struct.colour_cast = int(bool(colour_cast))
        eng = _ENGINES[key] = Engine(library, h, w, max_batch=max_batch, strategies=strategies, weights=weights,
                                     lib_struct=lib_struct)
    return eng


def _clahe_rgb_bytes(frame: np.ndarray, clip_limit, tile_grid_size, library):
    """RGB bytes -> Lab -> CLAHE on L -> RGB bytes (ES:299-306, SS:204-207) through a one-strategy engine."""
    s = _lib.Strategy()
    s.clahe, s.clip_limit = 1, float(clip_limit)
    s.tiles_x, s.tiles_y = (int(v) for v in tile_grid_size)
    eng = _engine(library, frame.shape[0], frame.shape[1], strategies=[s], colour_cast=False)
    return eng.run_host(frame[None], want_all=True)["out"][0, 0]


# ------------------------------------------------------------------ enhancement_strategies.py (float64 results)
class ESApi:
    """EnhancementStrategies of enhancement_strategies.py:13-508: same static methods, defaults and return types."""

    @staticmethod
    def guided_filter(I, p, r, eps):                                           # ES:16-46
        return prims.guided_filter(np.asarray(I, np.float64), np.asarray(p, np.float64), r, eps).cpu().numpy()

    @staticmethod
    def estimate_atmospheric_light(image, min_size=1):                         # ES:77-144 -> (H, W, 3) tile
        if min_size != 1:
            raise ValueError("estimate_atmospheric_light: only min_size=1 (the value every caller passes) is built")
        a = _check_image(image)
        light = prims.atmospheric_light(_float_tensor(a)).cpu().numpy()
        if a.dtype in (np.float32, np.float64):
            light = light.astype(a.dtype)
        return np.tile(light.reshape(1, 1, 3), (a.shape[0], a.shape[1], 1))     # ES:142

    @staticmethod
    def compute_Q(block):                                                      # ES:147-188
        return prims.region_score(_float_tensor(block))

    @staticmethod
    def get_brightest_pixel(block):                                            # ES:191-206
        return prims.brightest_pixel(_float_tensor(block)).cpu().numpy()

    @staticmethod
    def _light3(A):
        A = np.asarray(A)
        if A.ndim == 3:
            h, w = A.shape[:2]
            probe = A[[0, h // 2, h - 1], [0, w // 2, w - 1]]
            if not np.all(probe == probe[0]):
                raise ValueError("atmospheric light must be one colour (the tile estimate_atmospheric_light returns)")
            return A[0, 0]
        return A.reshape(-1)[:3]

    @staticmethod
    def estimate_transmission(img, A, omega=0.95, r=15, eps=0.001):            # ES:209-234 -> float64 (H, W)
        return prims.transmission(_float_tensor(img), ESApi._light3(A), omega, r, eps, 1e-10, False).cpu().numpy()

    @staticmethod
    def recover_image(img, t, A):                                              # ES:237-249 -> float64
        return prims.recover(_float_tensor(img), np.asarray(t, np.float64), ESApi._light3(A), torch.float64).cpu().numpy()

    @staticmethod
    def color_enhancement(img, L_low=15, L_high=95):                           # ES:252-273 -> dtype of img
        return prims.percentile_stretch(_float_tensor(img), L_low, L_high, 1e-10).cpu().numpy()

    @staticmethod
    def gamma_correction(img, gamma=1.2):                                      # ES:276-285
        return prims.gamma(_float_tensor(img), 1.0 / gamma, True).cpu().numpy()

    @staticmethod
    def clahe_enhancement(img, clip_limit=2.0, tile_grid_size=(8, 8)):         # ES:288-307 -> float64, k / 255
        q = prims.quantise(_float_tensor(img)).cpu().numpy() if as_frame(img) is None else as_frame(img)
        return _clahe_rgb_bytes(q, clip_limit, tile_grid_size, "es").astype(np.float64) / 255.0

    @staticmethod
    def histogram_equalization(img):                                           # ES:331-345 -> float64, k / 255
        f = as_frame(img)
        q = prims.quantise(_float_tensor(img)) if f is None else prims.as_device(f)
        return prims.equalize_hist(q).cpu().numpy().astype(np.float64) / 255.0

    # ---- apply_* (ES:349-474): the engine for byte-valued images, the primitives otherwise
    @staticmethod
    def _apply(img, name, params):
        frame = as_frame(img)
        if frame is not None:
            eng = _engine("es", frame.shape[0], frame.shape[1], strategies=[_es_strategy(name, params)])
            return eng.run_host(frame[None], want_all=True)["out"][0, 0].astype(np.float64) / 255.0
        d = _ES_DEFAULTS[name]
        g = lambda k: params.get(k, d[k])                                      # noqa: E731
        t = _float_tensor(img)
        if d["kind"] == "dehaze":
            light = prims.atmospheric_light(t)
            tr = prims.transmission(t, light, g("omega"), int(g("guided_radius")), 1e-3, 1e-10, False)
            cur = prims.recover(t, tr, light, torch.float64)
        elif d["kind"] == "clahe":
            q = prims.quantise(t).cpu().numpy()
            cur = prims.as_device(_clahe_rgb_bytes(q, g("clip_limit"), g("tile_grid_size"), "es")).to(torch.float64) / 255.0
        else:
            cur = prims.equalize_hist(prims.quantise(t)).to(torch.float64) / 255.0
        cur = prims.percentile_stretch(cur.contiguous(), g("L_low"), g("L_high"), 1e-10)
        if params.get("apply_gamma", False):
            cur = prims.gamma(cur, 1.0 / params.get("gamma", 1.2), True)
        return cur.cpu().numpy()

    @staticmethod
    def apply_strong_dehazing(img, params):
        return ESApi._apply(img, "strong_dehazing", params)

    @staticmethod
    def apply_medium_dehazing(img, params):
        return ESApi._apply(img, "medium_dehazing", params)

    @staticmethod
    def apply_light_enhancement(img, params):
        return ESApi._apply(img, "light_enhancement", params)

    @staticmethod
    def apply_clahe_enhancement(img, params):
        return ESApi._apply(img, "clahe_enhancement", params)

    @staticmethod
    def apply_histogram_equalization(img, params):
        return ESApi._apply(img, "histogram_equalization", params)

    @staticmethod
    def apply_strategy(img, strategy_name, params):
        """ValueError for an unknown name (ES:500-501); any failure inside the strategy prints and returns the input
        image (ES:506-508)."""
        if strategy_name not in _ES_DEFAULTS:
            raise ValueError(f"未知策略: {strategy_name}")
        try:
            return ESApi._apply(img, strategy_name, params)
        except Exception as e:                                                  # noqa: BLE001  (the reference's behaviour)
            print(f"策略 {strategy_name} 執行失敗: {e}")
            return img


es_apply_strategy = ESApi.apply_strategy


def enhance_and_select(img, strategies: dict | None = None, weights: dict | None = None):
    """main.py:116-155 for one byte-valued image: every strategy of `strategies` (default Config.STRATEGIES),
    comprehensive_assessment of each with `weights` (default Config.QUALITY_WEIGHTS), first maximum.
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


# ------------------------------------------------------------------ six_stadigy.py (float32 results)
_SS_CHAINS = (          # six_stadigy.py:230-285 as data: (op, args...)
    (("dehaze", 0.3, 20, 5e-1), ("stretch", 5, 98), ("clahe", 3.0), ("gamma", 1.5)),
    (("dehaze", 0.5, 15, 5e-1), ("stretch", 15, 95), ("clahe", 2.0)),
    (("dehaze", 0.7, 10, 1e-1), ("stretch", 20, 85), ("stretch", 2, 98)),
    (("clahe", 4.0), ("stretch", 10, 95), ("stretch", 3, 97), ("gamma", 1.3)),
    (("stretch", 2, 98), ("stretch", 15, 90), ("clahe", 1.5), ("gamma", 1.2)),
    (("stretch", 5, 98), ("clahe", 3.5), ("gamma", 1.4)),
)


class SSApi:
    """EnhancementStrategies of six_stadigy.py:22-285: same methods, defaults and return types (float32 images)."""

    guided_filter = staticmethod(ESApi.guided_filter)                          # SS:25-46
    compute_Q = staticmethod(ESApi.compute_Q)                                  # SS:115-157
    get_brightest_pixel = staticmethod(ESApi.get_brightest_pixel)              # SS:159-165

    @staticmethod
    def estimate_atmospheric_light(image, min_size=1):                         # SS:48-113 -> (3,)
        if min_size != 1:
            raise ValueError("estimate_atmospheric_light: only min_size=1 (the value every caller passes) is built")
        a = _check_image(image)
        light = prims.atmospheric_light(_float_tensor(a)).cpu().numpy()
        return light.astype(a.dtype) if a.dtype in (np.float32, np.float64) else light

    @staticmethod
    def estimate_transmission(img, A, omega, guided_radius, eps):              # SS:167-180 -> float64 (H, W)
        return prims.transmission(_float_tensor(img), ESApi._light3(A), omega, guided_radius, eps, 1e-6, True).cpu().numpy()

    @staticmethod
    def restore_image(img, A, t):                                              # SS:182-188 -> dtype of img
        x = _float_tensor(img)
        return prims.recover(x, np.asarray(t, np.float64), ESApi._light3(A), x.dtype).cpu().numpy()

    @staticmethod
    def enhance_contrast(img, L_low, L_high):                                  # SS:190-199
        return prims.percentile_stretch(_float_tensor(img), L_low, L_high, 1e-6).cpu().numpy()

    @staticmethod
    def apply_clahe(img, clip_limit=2.0):                                      # SS:201-208 -> float32, k / 255
        f = as_frame(img)
        q = prims.quantise(_float_tensor(img)).cpu().numpy() if f is None else f
        return _clahe_rgb_bytes(q, clip_limit, (8, 8), "six").astype(np.float32) / np.float32(255.0)

    @staticmethod
    def white_balance(img, percentile=5):                                      # SS:210-219
        return prims.percentile_stretch(_float_tensor(img), percentile, 100 - percentile, 1e-6).cpu().numpy()

    @staticmethod
    def gamma_correction(img, gamma=1.2):                                      # SS:221-224: img ** gamma, no clip
        return prims.gamma(_float_tensor(img), gamma, False).cpu().numpy()

    @staticmethod
    def _strategy(index, img):
        frame = as_frame(img)
        if frame is not None:                                                   # the engine, without the colour cast
            _, st = _lib.preset(_lib.LIB_SIX)
            eng = _engine("six", frame.shape[0], frame.shape[1], strategies=[st[index]], colour_cast=False)
            return eng.run_host(frame[None], want_all=True)["out"][0, 0].astype(np.float32) / np.float32(255.0)
        cur = _float_tensor(img)
        light = None
        for step in _SS_CHAINS[index]:
            if step[0] == "dehaze":
                light = prims.atmospheric_light(cur)
                tr = prims.transmission(cur, light, step[1], step[2], step[3], 1e-6, True)
                cur = prims.recover(cur, tr, light, cur.dtype)
            elif step[0] == "stretch":
                cur = prims.percentile_stretch(cur, step[1], step[2], 1e-6)
            elif step[0] == "clahe":
                q = prims.quantise(cur).cpu().numpy()
                cur = prims.as_device(_clahe_rgb_bytes(q, step[1], (8, 8), "six")).to(torch.float32) / 255.0
            else:
                cur = prims.gamma(cur, step[1], False)
        return cur.cpu().numpy()

    @classmethod
    def strategy1_strong_dehazing(cls, img):
        return cls._strategy(0, img)

    @classmethod
    def strategy2_medium_dehazing(cls, img):
        return cls._strategy(1, img)

    @classmethod
    def strategy3_light_dehazing(cls, img):
        return cls._strategy(2, img)

    @classmethod
    def strategy4_clahe_enhancement(cls, img):
        return cls._strategy(3, img)

    @classmethod
    def strategy5_white_balance(cls, img):
        return cls._strategy(4, img)

    @classmethod
    def strategy6_histogram_eq(cls, img):
        return cls._strategy(5, img)


def detect_image_type(img) -> str:
    """six_stadigy.py:292-302: channel means only (three sums on the device, nothing else runs)."""
    a = _check_image(img)
    if a.dtype == np.uint8:
        eng = _engine("six", a.shape[0], a.shape[1])
        return IMAGE_TYPES[int(eng.image_types(torch.from_numpy(np.ascontiguousarray(a))[None].to(eng.device))[0])]
    r, g, b = (float(v) for v in prims.channel_means(_float_tensor(a)).cpu())
    if g > r and g > b and (g - r) > 0.05:
        return "greenish"
    if b > r and b > g and (b - r) > 0.05:
        return "bluish"
    return "normal"


def color_correction(img, image_type):
    """six_stadigy.py:305-323: G (greenish) or B (bluish) x 0.85, clipped; `normal` returns the same array."""
    if image_type not in ("greenish", "bluish"):
        return img
    scale = (1.0, 0.85, 1.0) if image_type == "greenish" else (1.0, 1.0, 0.85)
    return prims.scale_clip(_float_tensor(img), scale).cpu().numpy()


def six_strategies(img, with_scores: bool = False):
    """six_stadigy.py:409-431 for one byte-valued image: (image_type, {strategy name: float32 image}) and, on
    request, the comprehensive_assessment rows of the six outputs."""
    frame = _to_frame(img)
    eng = _engine("six", frame.shape[0], frame.shape[1])
    res = eng.run_host(frame[None], want_all=True)
    kind = IMAGE_TYPES[int(res["type"][0])]
    outs = {n: res["out"][0, i].astype(np.float32) / np.float32(255.0) for i, n in enumerate(SS_NAMES)}
    if with_scores:
        return kind, outs, {n: res["scores"][0, i].copy() for i, n in enumerate(SS_NAMES)}
    return kind, outs


def process_all_images_all_strategies(input_folder, output_base_folder, batch: int = 8, workers: int = 4,
                                      write_csv: bool = True):
    """six_stadigy.py:326-470: every image of `input_folder` through colour cast + the six strategies, written as
    <stem>_<strategy name>.png under `output_base_folder`, with the reference's processing_log.csv
    (filename, image_type, strategy, strategy_desc, status, output_path, processing_time).

    Streaming: files are decoded ahead on worker threads into pinned slots, `batch` frames of one size at a time go
    through the engine's host entry point, and the 6 x batch outputs are encoded on worker threads while the next
    batch runs (frame_io.py).  A batch that fails logs the reference's 'failed' rows and the loop goes on.
    Returns the log rows."""
    from .frame_io import FrameReader, FrameWriter, list_images
    src, dst = Path(input_folder), Path(output_base_folder)
    dst.mkdir(parents=True, exist_ok=True)
    files = list_images(src)
    if not files:
        print(f"❌ 在 {input_folder} 中找不到影像檔案！")
        return []
    log = []
    reader, writer = FrameReader(files, batch=batch, workers=workers), FrameWriter(workers=workers)
    pending = []                                            # (row dict, future) of writes in flight
    try:
        for paths, frames in reader:
            t0 = time.time()
            n, h, w = frames.shape[:3]
            try:
                eng = _engine("six", h, w, max_batch=batch)
                out = writer.slot((batch, len(SS_NAMES), h, w, 3))[:n]
                res = eng.run_host(frames, want_all=True, out_all=out)
            except Exception as e:                          # noqa: BLE001  (six_stadigy.py:447-470: log and go on)
                for p in paths:
                    for name, desc in zip(SS_NAMES, SS_DESCRIPTIONS):
                        log.append(dict(filename=p.name, image_type="", strategy=name, strategy_desc=desc, status="failed",
                                        output_path=f"Error: {str(e)[:50]}", processing_time="N/A"))
                continue
            dt = f"{(time.time() - t0) / n:.2f}s"
            for i, p in enumerate(paths):
                kind = IMAGE_TYPES[int(res["type"][i])]
                for s, (name, desc) in enumerate(zip(SS_NAMES, SS_DESCRIPTIONS)):
                    target = dst / f"{p.stem}_{name}.png"
                    row = dict(filename=p.name, image_type=kind, strategy=name, strategy_desc=desc, status="success",
                               output_path=str(target), processing_time=dt)
                    log.append(row)
                    pending.append((row, writer.write(target, out[i, s])))
        for p in reader.failed:
            print(f"  無法讀取: {p.name}")
    finally:
        for row, fut in pending:
            try:
                if not fut.result():
                    row["status"], row["output_path"] = "failed", "Error: imwrite"
            except Exception as e:                          # noqa: BLE001
                row["status"], row["output_path"], row["processing_time"] = "failed", f"Error: {str(e)[:50]}", "N/A"
        writer.close()
        reader.close()
    if write_csv:
        import pandas as pd
        pd.DataFrame(log).to_csv(dst / "processing_log.csv", index=False, encoding="utf-8-sig")   # SS:450-459
    return log


def build_dataset_rows(image_files, strategies: dict, weights: dict, strategy_folder=None, batch: int = 8,
                       workers: int = 4, with_features: bool = True):
    """The loop body of SelfSupervisedSystem.build_dataset (main.py:88-183), batched and streamed: for every readable
    image of at least 10 x 10 pixels the 79 features of the raw image (main.py:116), every strategy's total score with
    `weights`, the first maximum, and the selected image written as <stem>_<best name>.png (main.py:150-155).
    Features, scores and the selection are produced on the device from the same resident frames.
    Returns (dataset entries like main.py:158-164, csv rows like main.py:167-173)."""
    from .features import extract_all_features_batch
    from .frame_io import FrameReader, FrameWriter
    keys = list(strategies)
    for k in keys:
        if k not in _ES_DEFAULTS:
            raise ValueError(f"未知策略: {k}")
    names = [strategies[k].get("name", k) for k in keys]
    strat = [_es_strategy(k, strategies[k]) for k in keys]
    dataset, rows = [], []
    reader, writer = FrameReader(image_files, batch=batch, workers=workers), FrameWriter(workers=workers)
    try:
        for paths, frames in reader:
            n, h, w = frames.shape[:3]
            if h < 10 or w < 10:                             # main.py:111-113
                for p in paths:
                    print(f"警告: {p.name} 尺寸過小，跳過")
                continue
            try:
                eng = _engine("es", h, w, strategies=strat, weights=weights, max_batch=batch)
                out = writer.slot((batch, h, w, 3))[:n] if strategy_folder else None
                res = eng.run_host(frames, want_all=False, want_best=strategy_folder is not None, out_best=out)
                feats = extract_all_features_batch(frames).cpu().numpy() if with_features else [None] * n
            except Exception as e:                          # noqa: BLE001  (main.py:181-183)
                for p in paths:
                    print(f"處理 {p.name} 時發生錯誤: {e}")
                continue
            for i, p in enumerate(paths):
                scores = {nm: float(res["scores"][i, j, 0]) for j, nm in enumerate(names)}
                best = names[int(res["best"][i])]
                if strategy_folder:
                    writer.write(Path(strategy_folder) / f"{p.stem}_{best}.png", out[i])
                dataset.append({"filename": p.name, "features": feats[i], "best_strategy": best,
                                "best_score": scores[best], "all_scores": scores})
                row = {"filename": p.name, "best_strategy": best, "best_score": scores[best]}
                row.update(scores)
                rows.append(row)
        for p in reader.failed:
            print(f"警告: 無法讀取 {p.name}")
    finally:
        writer.close()
        reader.close()
    return dataset, rows


def close_engines():
    while _ENGINES:
        _ENGINES.popitem()[1].close()
