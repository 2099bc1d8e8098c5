"""PyTorch operators over CUDA tensors: torch.ops.uwie.* (SURVEY.md §8 b-2).

Importing this module registers the operators (CUDA dispatch key only: there is no CPU kernel, calling
them with CPU tensors fails in the dispatcher).  They are thin: every one hands raw device pointers and the
current stream to the C ABI of libuwie_b200.so through ctypes.

    uwie::enhance_score_p6(Tensor frames) -> (Tensor out, Tensor scores, Tensor best, Tensor type)
        uint8 [B,H,W,3] -> uint8 [B,6,H,W,3], float64 [B,6,9], int32 [B], int32 [B]      six_stadigy.py:409-431 + QA
    uwie::enhance_score_p5(Tensor frames) -> (Tensor best_out, Tensor scores, Tensor best)
        uint8 [B,H,W,3] -> uint8 [B,H,W,3], float64 [B,5,9], int32 [B]                  main.py:122-155
    uwie::quality_scores(Tensor img, float[] weights) -> Tensor                          quality_assessment.py:215-286
    uwie::features79(Tensor frames) -> Tensor                                            feature_extraction.py:249-297
    uwie::vgg_enhance(Tensor img, Tensor l_low, Tensor l_high, Tensor? omega, Tensor? gamma) -> Tensor   vgg_16_UIE.py:32-55
    uwie::vgg_features(Tensor frames) -> Tensor                                          vgg_16_UIE.py:435-466
    uwie::vgg_preprocess(Tensor frames, int size) -> Tensor                              use_trained_model.py:39-46
Per-primitive operators (csrc/k_prims.cu; float images are [H,W,3] float32 / float64 and need not be byte-valued):
    uwie::clahe_lab(Tensor planes, float clip_limit, int tiles_x, int tiles_y) -> Tensor     uint8 [B,H,W]      ES:302-303, SS:205-206
    uwie::equalize_hist(Tensor frames) -> Tensor                                           uint8 [B,H,W,3]    ES:339-343
    uwie::transmission_head(Tensor frames, Tensor light, float omega, float guard, bool clip) -> Tensor   uint8 -> float32 [B,H,W]  SS:170-174
    uwie::atmospheric_light(Tensor img) -> Tensor                                          (3,)               ES:77-206, SS:48-165
    uwie::transmission(Tensor img, Tensor light, float omega, int r, float eps, float guard, bool clip_t0) -> Tensor   float64 [H,W]  ES:209-234, SS:167-180
    uwie::recover(Tensor img, Tensor t, Tensor light, bool f64_out) -> Tensor             ES:237-249, SS:182-188
    uwie::percentile_stretch(Tensor img, float lo, float hi, float guard) -> Tensor        ES:252-273, SS:190-219
    uwie::gamma(Tensor img, float exponent, bool clip) -> Tensor                           ES:276-285, SS:221-224
"""
from __future__ import annotations

import torch

from . import _lib
from .engine import QA_KEYS, Engine

_lib.load()          # fail at import time, loudly, if the native library is missing

_LIB = torch.library.Library("uwie", "DEF")
_LIB.define("enhance_score_p6(Tensor frames) -> (Tensor, Tensor, Tensor, Tensor)")
_LIB.define("enhance_score_p5(Tensor frames) -> (Tensor, Tensor, Tensor)")
_LIB.define("quality_scores(Tensor img, float[] weights) -> Tensor")
_LIB.define("features79(Tensor frames) -> Tensor")
_LIB.define("vgg_enhance(Tensor img, Tensor l_low, Tensor l_high, Tensor? omega, Tensor? gamma) -> Tensor")
_LIB.define("vgg_features(Tensor frames) -> Tensor")
_LIB.define("vgg_preprocess(Tensor frames, int size) -> Tensor")

_LIB.define("clahe_lab(Tensor planes, float clip_limit, int tiles_x, int tiles_y) -> Tensor")
_LIB.define("equalize_hist(Tensor frames) -> Tensor")
_LIB.define("transmission_head(Tensor frames, Tensor light, float omega, float guard, bool clip) -> Tensor")
_LIB.define("atmospheric_light(Tensor img) -> Tensor")
_LIB.define("transmission(Tensor img, Tensor light, float omega, int r, float eps, float guard, bool clip_t0) -> Tensor")
_LIB.define("recover(Tensor img, Tensor t, Tensor light, bool f64_out) -> Tensor")
_LIB.define("percentile_stretch(Tensor img, float lo, float hi, float guard) -> Tensor")
_LIB.define("gamma(Tensor img, float exponent, bool clip) -> Tensor")

_ENGINES: dict = {}
_CHUNK = 16          # frames per engine call: fewer, fuller launches (bench.py --chunk)


def _engine(library, frames):
    key = (library, frames.shape[1], frames.shape[2], frames.device.index)
    eng = _ENGINES.get(key)
    if eng is None:
        eng = _ENGINES[key] = Engine(library, frames.shape[1], frames.shape[2], max_batch=_CHUNK, device=str(frames.device))
    return eng


def _check_frames(frames):
    if frames.dtype != torch.uint8 or frames.dim() != 4 or frames.shape[-1] != 3:
        raise ValueError(f"frames must be uint8 [B,H,W,3], got {tuple(frames.shape)} {frames.dtype}")
    return frames.contiguous()


def _run(library, frames, want_all, want_best):
    frames = _check_frames(frames)
    eng = _engine(library, frames)
    B, H, W, _ = frames.shape
    dev = frames.device
    out_all = torch.empty((B, eng.S, H, W, 3), dtype=torch.uint8, device=dev) if want_all else None
    out_best = torch.empty((B, H, W, 3), dtype=torch.uint8, device=dev) if want_best else None
    scores = torch.empty((B, eng.S, 9), dtype=torch.float64, device=dev)
    best = torch.empty((B,), dtype=torch.int32, device=dev)
    types = torch.empty((B,), dtype=torch.int32, device=dev)
    for c0 in range(0, B, _CHUNK):
        c1 = min(B, c0 + _CHUNK)
        eng.run(frames[c0:c1], want_all=want_all, want_best=want_best,
                out_all=None if out_all is None else out_all[c0:c1],
                out_best=None if out_best is None else out_best[c0:c1],
                scores=scores[c0:c1], best=best[c0:c1], types=types[c0:c1])
    eng.check()     # device-side failures (percentile selection overflow) must not pass as plausible outputs
    return out_all, out_best, scores, best, types


def _p6(frames):
    out_all, _, scores, best, types = _run("six", frames, True, False)
    return out_all, scores, best, types


def _p5(frames):
    _, out_best, scores, best, _ = _run("es", frames, False, True)
    return out_best, scores, best


def _quality(img, weights):
    from .quality import quality_scores_batch
    if len(weights) != 8:
        raise ValueError("weights: 8 values in QA key order " + ", ".join(QA_KEYS))
    return quality_scores_batch(img, dict(zip(QA_KEYS, weights)), device=img.device)


def _features79(frames):
    from .features import extract_all_features_batch
    return extract_all_features_batch(_check_frames(frames), device=frames.device)


def _vgg_enhance(img, l_low, l_high, omega, gamma):
    from .vgg import DifferentiableEnhancement
    params = {"L_low": l_low, "L_high": l_high}
    if omega is not None:
        params["omega"] = omega
    if gamma is not None:
        params["gamma"] = gamma
    return DifferentiableEnhancement().forward(img, params)


def _vgg_features(frames):
    from .vgg import extract_all_features
    return extract_all_features(_check_frames(frames), device=frames.device)


def _vgg_preprocess(frames, size):
    from .vgg import preprocess_for_vgg
    return preprocess_for_vgg(frames, size, device=frames.device)


def _prim(name):
    from . import prims
    return getattr(prims, name)


_LIB.impl("clahe_lab", lambda planes, clip, tx, ty: _prim("clahe_lab")(planes, clip, (tx, ty)), "CUDA")
_LIB.impl("equalize_hist", lambda frames: _prim("equalize_hist")(frames), "CUDA")
_LIB.impl("transmission_head", lambda frames, light, omega, guard, clip: _prim("transmission_head_u8")(frames, light, omega, guard, clip), "CUDA")
_LIB.impl("atmospheric_light", lambda img: _prim("atmospheric_light")(img), "CUDA")
_LIB.impl("transmission", lambda img, light, omega, r, eps, guard, clip_t0: _prim("transmission")(img, light, omega, r, eps, guard, clip_t0), "CUDA")
_LIB.impl("recover", lambda img, t, light, f64_out: _prim("recover")(img, t, light, torch.float64 if f64_out else torch.float32), "CUDA")
_LIB.impl("percentile_stretch", lambda img, lo, hi, guard: _prim("percentile_stretch")(img, lo, hi, guard), "CUDA")
_LIB.impl("gamma", lambda img, exponent, clip: _prim("gamma")(img, exponent, clip), "CUDA")
_LIB.impl("enhance_score_p6", _p6, "CUDA")
_LIB.impl("enhance_score_p5", _p5, "CUDA")
_LIB.impl("quality_scores", _quality, "CUDA")
_LIB.impl("features79", _features79, "CUDA")
_LIB.impl("vgg_enhance", _vgg_enhance, "CUDA")
_LIB.impl("vgg_features", _vgg_features, "CUDA")
_LIB.impl("vgg_preprocess", _vgg_preprocess, "CUDA")


def close_engines():
    while _ENGINES:
        _ENGINES.popitem()[1].close()
