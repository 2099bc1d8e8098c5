"""Host-side mirror of feature_extraction.py's FeatureExtractor for the 79-feature vector.

All arithmetic runs in libuwie_b200.so (uwie_features79); PyTorch provides device memory and the
current stream.  Group order and sizes follow FE:249-297.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib

GROUPS = (("color", 0, 35), ("texture", 35, 57), ("frequency", 57, 62), ("edge", 62, 69), ("quality", 69, 79))


def _as_frames(img, device):
    """uint8 frames, or the reference's float images in [0, 1] (truncated like (img * 255).astype(uint8),
    FE:30) -> contiguous CUDA uint8 [B, H, W, 3]."""
    t = img if isinstance(img, torch.Tensor) else torch.as_tensor(np.ascontiguousarray(img))
    if t.dim() == 3:
        t = t[None]
    if t.dim() != 4 or t.shape[-1] != 3:
        raise ValueError(f"expected (H, W, 3) or (B, H, W, 3), got {tuple(t.shape)}")
    t = t.to(device)
    if t.dtype != torch.uint8:
        if not t.is_floating_point():
            raise TypeError(f"unsupported dtype {t.dtype}")
        t = (t * 255).to(torch.uint8)          # dtype conversion only: every feature is computed natively
    return t.contiguous()


def extract_all_features_batch(img, device="cuda:0") -> torch.Tensor:
    """float64 [B, 79] on the device."""
    lib = _lib.load()
    if not torch.cuda.is_available():
        raise RuntimeError("uwie: no CUDA device — the feature path has no CPU fallback")
    frames = _as_frames(img, device)
    B, H, W, _ = frames.shape
    need = lib.uwie_features79_workspace_bytes(B, H, W)
    if need == 0:
        _lib.check(_lib.ERR_ARG)
    ws = _lib.scratch(need, frames.device)
    feat = torch.empty((B, _lib.FEATURES), dtype=torch.float64, device=frames.device)
    stream = torch.cuda.current_stream(frames.device)
    with torch.cuda.device(frames.device):
        _lib.check(lib.uwie_features79(C.c_void_p(frames.data_ptr()), B, H, W, C.c_void_p(feat.data_ptr()),
                                       C.c_void_p(ws.data_ptr()), need, C.c_void_p(stream.cuda_stream)))
    ws.record_stream(stream)
    frames.record_stream(stream)
    return feat


class FeatureExtractor:
    """feature_extraction.py:13-297 — same static methods, NumPy float64 vectors out."""

    @staticmethod
    def extract_all_features(img) -> np.ndarray:
        return extract_all_features_batch(img)[0].cpu().numpy()

    @staticmethod
    def _group(img, name):
        lo, hi = next((a, b) for n, a, b in GROUPS if n == name)
        return extract_all_features_batch(img)[0, lo:hi].cpu().numpy()

    @staticmethod
    def extract_color_features(img) -> np.ndarray:
        return FeatureExtractor._group(img, "color")

    @staticmethod
    def extract_texture_features(img) -> np.ndarray:
        return FeatureExtractor._group(img, "texture")

    @staticmethod
    def extract_frequency_features(img) -> np.ndarray:
        return FeatureExtractor._group(img, "frequency")

    @staticmethod
    def extract_edge_features(img) -> np.ndarray:
        return FeatureExtractor._group(img, "edge")

    @staticmethod
    def extract_quality_features(img) -> np.ndarray:
        return FeatureExtractor._group(img, "quality")
