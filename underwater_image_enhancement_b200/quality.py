"""Host-side mirror of quality_assessment.py's QualityAssessment for images that did not come out of an
Engine (the fused path scores its own outputs).  Arithmetic: libuwie_b200.so (uwie_quality_scores)."""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from .engine import QA_KEYS

DEFAULT_WEIGHTS = {"contrast": 0.20, "sharpness": 0.20, "entropy": 0.15, "saturation": 0.15, "brightness": 0.10,
                   "edge_density": 0.10, "colorfulness": 0.05, "naturalness": 0.05}      # QA:229-238
_DTYPES = {torch.uint8: 0, torch.float32: 1, torch.float64: 2}


def quality_scores_batch(img, weights=None, device="cuda:0") -> torch.Tensor:
    """img: (H, W, 3) or (B, H, W, 3), uint8 (meaning bytes / 255), float32 or float64 in [0, 1], NumPy or
    torch.  Returns float64 [B, 9] on the device: total, then the eight sub-scores in QA key order."""
    lib = _lib.load()
    if not torch.cuda.is_available():
        raise RuntimeError("uwie: no CUDA device — the scoring path has no CPU fallback")
    t = img if isinstance(img, torch.Tensor) else torch.as_tensor(np.ascontiguousarray(img))
    if t.dim() == 3:
        t = t[None]
    if t.dim() != 4 or t.shape[-1] != 3:
        raise ValueError(f"expected (H, W, 3) or (B, H, W, 3), got {tuple(t.shape)}")
    if t.dtype not in _DTYPES:
        raise TypeError(f"unsupported dtype {t.dtype}")
    t = t.to(device).contiguous()
    B, H, W, _ = t.shape
    w = DEFAULT_WEIGHTS if weights is None else weights
    warr = (C.c_double * 8)(*[float(w.get(k, 0)) for k in QA_KEYS])                      # QA:284 weights.get(k, 0)
    need = lib.uwie_quality_workspace_bytes(B, H, W)
    ws = _lib.scratch(need, t.device)
    scores = torch.empty((B, _lib.SCORE_COLS), dtype=torch.float64, device=t.device)
    stream = torch.cuda.current_stream(t.device)
    with torch.cuda.device(t.device):
        _lib.check(lib.uwie_quality_scores(C.c_void_p(t.data_ptr()), _DTYPES[t.dtype], B, H, W, warr,
                                           C.c_void_p(scores.data_ptr()), C.c_void_p(ws.data_ptr()), need,
                                           C.c_void_p(stream.cuda_stream)))
    ws.record_stream(stream)
    t.record_stream(stream)
    return scores


class QualityAssessment:
    """quality_assessment.py:11-286 — same static methods and return types."""

    @staticmethod
    def comprehensive_assessment(img, weights=None):
        row = quality_scores_batch(img, weights)[0].cpu().numpy()
        return float(row[0]), {k: float(row[1 + i]) for i, k in enumerate(QA_KEYS)}

    @staticmethod
    def _one(img, key):
        return QualityAssessment.comprehensive_assessment(img)[1][key]


for _k in QA_KEYS:
    setattr(QualityAssessment, "assess_" + _k, staticmethod(lambda img, _k=_k: QualityAssessment._one(img, _k)))
