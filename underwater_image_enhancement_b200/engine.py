"""Python face of the native engine: torch CUDA tensors in, torch CUDA tensors out.

PyTorch is used for device memory and streams only; all arithmetic is in libuwie_b200.so.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib

SS_NAMES = ("strong_dehazing", "medium_dehazing", "light_dehazing", "clahe_enhancement",
            "white_balance", "histogram_eq")                       # six_stadigy.py:373-380
ES_NAMES = ("strong_dehazing", "medium_dehazing", "light_enhancement", "clahe_enhancement",
            "histogram_equalization")                              # config.py:29-74
QA_KEYS = ("contrast", "sharpness", "entropy", "saturation", "brightness", "edge_density",
           "colorfulness", "naturalness")                          # quality_assessment.py:243-281
IMAGE_TYPES = ("normal", "greenish", "bluish")                     # six_stadigy.py:292-302


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


class Engine:
    """One engine = one strategy library x one frame size x one maximum batch on one device."""

    def __init__(self, library="six", height=1080, width=1920, max_batch=8, strategies=None,
                 weights=None, device="cuda:0", lib_struct=None):
        self._h = None
        lib = _lib.load()
        if not torch.cuda.is_available():
            raise RuntimeError("uwie: no CUDA device — the hot path has no CPU fallback")
        self.device = torch.device(device)
        lib_id = {"six": _lib.LIB_SIX, "ss": _lib.LIB_SIX, "es": _lib.LIB_ES}[library] if isinstance(library, str) else int(library)
        plib, pst = _lib.preset(lib_id)
        if lib_struct is not None:
            plib = lib_struct
        if strategies is not None:
            pst = list(strategies)
        if weights is not None:
            for i, k in enumerate(QA_KEYS):
                plib.weights[i] = float(weights.get(k, 0)) if isinstance(weights, dict) else float(weights[i])
        self.lib_struct, self.strategies = plib, pst
        self.S, self.H, self.W, self.max_batch = len(pst), int(height), int(width), int(max_batch)
        self.f64 = bool(plib.f64)
        arr = (_lib.Strategy * len(pst))(*pst)
        with torch.cuda.device(self.device):
            need = lib.uwie_workspace_bytes(C.byref(plib), arr, len(pst), self.max_batch, self.H, self.W)
            if need == 0:
                _lib.check(_lib.ERR_ARG)
            self.workspace = _lib.scratch(need, self.device)       # poisoned / guarded under the test hooks
            h = C.c_void_p()
            _lib.check(lib.uwie_engine_create(C.byref(h), C.byref(plib), arr, len(pst), self.max_batch,
                                              self.H, self.W, C.c_void_p(self.workspace.data_ptr()), need))
        self._h = h
        self._lib = lib

    def close(self):
        if self._h is not None:
            self._lib.uwie_engine_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ device path
    def run(self, frames: torch.Tensor, want_all=True, want_best=False, out_all=None, out_best=None,
            scores=None, best=None, types=None, light=None):
        """frames: uint8 CUDA tensor [B, H, W, 3], B <= max_batch.  Asynchronous on the current stream."""
        if frames.dtype != torch.uint8 or frames.dim() != 4 or frames.shape[1:] != (self.H, self.W, 3):
            raise ValueError(f"frames must be uint8 [B,{self.H},{self.W},3], got {tuple(frames.shape)} {frames.dtype}")
        if not frames.is_cuda:
            raise ValueError("frames must live on the CUDA device (use run_host for host arrays)")
        frames = frames.contiguous()
        B = frames.shape[0]
        dev = frames.device
        if want_all and out_all is None:
            out_all = torch.empty((B, self.S, self.H, self.W, 3), dtype=torch.uint8, device=dev)
        if want_best and out_best is None:
            out_best = torch.empty((B, self.H, self.W, 3), dtype=torch.uint8, device=dev)
        for t in (out_all, out_best, scores, best, types, light):
            if t is not None and not (t.is_cuda and t.is_contiguous()):
                raise ValueError("output tensors must be contiguous CUDA tensors")
        if scores is None:
            scores = torch.empty((B, self.S, _lib.SCORE_COLS), dtype=torch.float64, device=dev)
        if best is None:
            best = torch.empty((B,), dtype=torch.int32, device=dev)
        if types is None:
            types = torch.empty((B,), dtype=torch.int32, device=dev)
        if light is None:
            light = torch.empty((B, 3), dtype=torch.float32, device=dev)
        stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        with torch.cuda.device(dev):
            _lib.check(self._lib.uwie_enhance_score(self._h, _ptr(frames), B, _ptr(out_all), _ptr(out_best),
                                                    _ptr(scores), _ptr(best), _ptr(types), _ptr(light), stream))
        return {"out": out_all, "best_out": out_best, "scores": scores, "best": best, "type": types,
                "light": light}

    def image_types(self, frames: torch.Tensor) -> torch.Tensor:
        """detect_image_type (six_stadigy.py:292-302) of uint8 CUDA frames [B, H, W, 3] -> int32 [B]
        (0 normal / 1 greenish / 2 bluish); only the channel histograms and the colour-cast set-up run."""
        frames = frames.contiguous()
        B = frames.shape[0]
        types = torch.empty((B,), dtype=torch.int32, device=frames.device)
        stream = C.c_void_p(torch.cuda.current_stream(frames.device).cuda_stream)
        with torch.cuda.device(frames.device):
            _lib.check(self._lib.uwie_engine_image_types(self._h, _ptr(frames), B, _ptr(types), stream))
        return types

    def check(self):
        """Synchronise and raise if the device reported a failure (e.g. selection overflow)."""
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        with torch.cuda.device(self.device):
            _lib.check(self._lib.uwie_engine_check(self._h, stream))

    @property
    def last_launches(self) -> int:
        return int(self._lib.uwie_engine_last_launches(self._h))

    @property
    def graph_replays(self) -> int:
        """Calls served by replaying a captured CUDA graph of the kernel sequence (uwie_engine_graph_replays)."""
        return int(self._lib.uwie_engine_graph_replays(self._h))

    # ------------------------------------------------------------------ host path (C ABI with host buffers)
    def run_host(self, frames: np.ndarray, want_all=True, want_best=False, out_all=None, out_best=None):
        frames = np.ascontiguousarray(frames, dtype=np.uint8)
        if frames.ndim == 3:
            frames = frames[None]
        if frames.shape[1:] != (self.H, self.W, 3):
            raise ValueError(f"frames must be uint8 [n,{self.H},{self.W},3], got {frames.shape}")
        n = frames.shape[0]
        if want_all and out_all is None:
            out_all = np.empty((n, self.S, self.H, self.W, 3), np.uint8)
        if want_best and out_best is None:
            out_best = np.empty((n, self.H, self.W, 3), np.uint8)
        scores = np.empty((n, self.S, _lib.SCORE_COLS), np.float64)
        best = np.empty((n,), np.int32)
        types = np.empty((n,), np.int32)
        light = np.empty((n, 3), np.float32)

        def hp(a):
            return None if a is None else a.ctypes.data_as(C.c_void_p)

        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        with torch.cuda.device(self.device):
            self.host_prepare(n, out_all is not None, out_best is not None)
            _lib.check(self._lib.uwie_enhance_score_host(self._h, hp(frames), n, hp(out_all), hp(out_best),
                                                         hp(scores), hp(best), hp(types), hp(light), stream))
        return {"out": out_all, "best_out": out_best, "scores": scores, "best": best, "type": types,
                "light": light}

    def host_prepare(self, max_count: int, want_all: bool = True, want_best: bool = False):
        """Allocate what run_host needs for up to `max_count` frames per call (uwie_engine_host_prepare):
        the host entry point itself allocates nothing.  Idempotent; run_host calls it on demand."""
        key = (int(max_count), bool(want_all), bool(want_best))
        have = getattr(self, "_host_ready", (0, False, False))
        if have[0] >= key[0] and (have[1] or not key[1]) and (have[2] or not key[2]):
            return
        with torch.cuda.device(self.device):
            _lib.check(self._lib.uwie_engine_host_prepare(self._h, key[0], int(key[1]), int(key[2])))
        self._host_ready = (max(have[0], key[0]), have[1] or key[1], have[2] or key[2])

    # ------------------------------------------------------------------ per-stage device timing
    def profile(self, enable: bool = True):
        _lib.check(self._lib.uwie_engine_profile(self._h, 1 if enable else 0))

    def profile_read(self):
        """[(stage, ms_total, kernels_launched, runs)] since profile(True); synchronises."""
        cap = 64
        names = C.create_string_buffer(cap * _lib.STAGE_NAME_LEN)
        ms = (C.c_double * cap)()
        ker = (C.c_int * cap)()
        runs = (C.c_int * cap)()
        with torch.cuda.device(self.device):
            n = self._lib.uwie_engine_profile_read(self._h, names, ms, ker, runs, cap)
        if n < 0:
            _lib.check(n)
        out = []
        for i in range(n):
            raw = names.raw[i * _lib.STAGE_NAME_LEN:(i + 1) * _lib.STAGE_NAME_LEN]
            out.append((raw.split(b"\0", 1)[0].decode(), float(ms[i]), int(ker[i]), int(runs[i])))
        return out

    # ------------------------------------------------------------------ test hook
    def debug(self, name: str, dtype, shape=None) -> np.ndarray:
        nbytes = C.c_size_t()
        p = self._lib.uwie_engine_debug(self._h, name.encode(), C.byref(nbytes))
        if not p:
            raise KeyError(name)
        off = p - self.workspace.data_ptr()
        torch.cuda.synchronize(self.device)
        raw = self.workspace[off:off + nbytes.value].cpu().numpy()
        arr = raw.view(dtype)
        return arr.reshape(shape) if shape is not None else arr
