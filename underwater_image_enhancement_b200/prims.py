"""Per-primitive operators over torch CUDA tensors: thin ctypes faces of the uwie_prim_* C ABI.

Every function takes and returns tensors on the device; NumPy arrays are accepted and moved there.  Float
images are [H, W, 3] float32 / float64 and need not be byte-valued.  All arithmetic runs in libuwie_b200.so.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib

_DT = {torch.float32: 1, torch.float64: 2}


def _dev(device=None):
    if not torch.cuda.is_available():
        raise RuntimeError("uwie: no CUDA device — the hot path has no CPU fallback")
    return torch.device(device or "cuda:0")


def as_device(x, device=None) -> torch.Tensor:
    if isinstance(x, torch.Tensor):
        if not x.is_cuda:
            x = x.to(_dev(device))
        return x.contiguous()
    return torch.from_numpy(np.ascontiguousarray(x)).to(_dev(device))


def _stream(t):
    return C.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)


def _p(t):
    return C.c_void_p(t.data_ptr())


def _ws(nbytes, dev):
    return _lib.scratch(nbytes, dev)


def _float_image(img, device=None):
    t = as_device(img, device)
    if t.dtype not in _DT:
        raise ValueError(f"expected a float32 / float64 image, got {t.dtype}")
    if t.dim() != 3 or t.shape[2] != 3:
        raise ValueError(f"expected an (H, W, 3) image, got {tuple(t.shape)}")
    return t


# ------------------------------------------------------------------ bandwidth-bound isolated kernels (configs[2])
def transmission_head_u8(frames, light, omega, guard=1e-6, clip_t0=True, out=None):
    """uint8 [B,H,W,3], float32 light [B,3] -> float32 t0 [B,H,W] (SS:170-174 / ES:221-225)."""
    frames = as_device(frames)
    light = as_device(light).to(torch.float32).contiguous()
    B, H, W, _ = frames.shape
    if out is None:
        out = torch.empty((B, H, W), dtype=torch.float32, device=frames.device)
    with torch.cuda.device(frames.device):
        _lib.check(_lib.load().uwie_prim_transmission_head_u8(_p(frames), B, H, W, _p(light), float(omega), float(guard),
                                                              int(bool(clip_t0)), _p(out), _stream(frames)))
    return out


def clahe_lab(plane, clip_limit=2.0, tile_grid_size=(8, 8), out=None, workspace=None):
    """cv2.createCLAHE(clip_limit, tile_grid_size).apply on uint8 planes [B,H,W] (or [H,W])."""
    plane = as_device(plane)
    squeeze = plane.dim() == 2
    if squeeze:
        plane = plane[None]
    if plane.dtype != torch.uint8 or plane.dim() != 3:
        raise ValueError("clahe_lab takes uint8 planes [B, H, W]")
    B, H, W = plane.shape
    tx, ty = int(tile_grid_size[0]), int(tile_grid_size[1])
    lib = _lib.load()
    need = lib.uwie_prim_clahe_lab_workspace_bytes(B, tx, ty)
    if need == 0:
        raise ValueError("CLAHE tile grid must have between 1 and 256 tiles")
    if workspace is None or workspace.numel() < need:
        workspace = _ws(need, plane.device)
    if out is None:
        out = torch.empty_like(plane)
    with torch.cuda.device(plane.device):
        _lib.check(lib.uwie_prim_clahe_lab(_p(plane), B, H, W, float(clip_limit), tx, ty, _p(out), _p(workspace),
                                           workspace.numel(), _stream(plane)))
    return out[0] if squeeze else out


def equalize_hist(frames):
    """cv2.equalizeHist per channel of uint8 [B,H,W,3] (or [H,W,3])."""
    frames = as_device(frames)
    squeeze = frames.dim() == 3
    if squeeze:
        frames = frames[None]
    B, H, W, _ = frames.shape
    out = torch.empty_like(frames)
    ws = _ws(B * 3840, frames.device)
    with torch.cuda.device(frames.device):
        _lib.check(_lib.load().uwie_prim_equalize_hist(_p(frames), B, H, W, _p(out), _p(ws), ws.numel(), _stream(frames)))
    return out[0] if squeeze else out


# ------------------------------------------------------------------ float-image primitives
def quantise(img):
    """(img * 255).astype(np.uint8) (truncation)."""
    t = as_device(img)
    if t.dtype not in _DT:
        raise ValueError("quantise takes float32 / float64")
    out = torch.empty(t.shape, dtype=torch.uint8, device=t.device)
    with torch.cuda.device(t.device):
        _lib.check(_lib.load().uwie_prim_quantise(_p(t), _DT[t.dtype], t.numel(), _p(out), _stream(t)))
    return out


def atmospheric_light(img):
    """estimate_atmospheric_light (ES:77-206, SS:48-165) -> (3,) tensor of the image's dtype."""
    t = _float_image(img)
    H, W, _ = t.shape
    lib = _lib.load()
    ws = _ws(lib.uwie_prim_atmospheric_light_workspace_bytes(H, W), t.device)
    light = torch.empty(3, dtype=t.dtype, device=t.device)
    with torch.cuda.device(t.device):
        _lib.check(lib.uwie_prim_atmospheric_light(_p(t), _DT[t.dtype], H, W, _p(light), _p(ws), ws.numel(), _stream(t)))
    return light


def guided_filter(guide, src, r, eps, clip=False):
    """guided_filter(I, p, r, eps) in float64 (ES:16-46, SS:25-46)."""
    g = as_device(guide).to(torch.float64).contiguous()
    s = as_device(src).to(torch.float64).contiguous()
    if g.dim() != 2 or g.shape != s.shape:
        raise ValueError("guided_filter takes two (H, W) planes")
    H, W = g.shape
    lib = _lib.load()
    ws = _ws(lib.uwie_prim_guided_filter_workspace_bytes(H, W), g.device)
    out = torch.empty_like(g)
    with torch.cuda.device(g.device):
        _lib.check(lib.uwie_prim_guided_filter(_p(g), _p(s), H, W, int(r), float(eps), int(bool(clip)), _p(out), _p(ws),
                                               ws.numel(), _stream(g)))
    return out


def transmission(img, light, omega, r, eps, guard, clip_t0):
    """estimate_transmission -> float64 (H, W) in [.1, 1] (ES:209-234 with guard 1e-10, clip_t0 False; SS:167-180 with
    guard 1e-6, clip_t0 True)."""
    t = _float_image(img)
    H, W, _ = t.shape
    light = as_device(light).to(t.dtype).reshape(-1)[:3].contiguous()
    lib = _lib.load()
    ws = _ws(lib.uwie_prim_transmission_workspace_bytes(H, W), t.device)
    out = torch.empty((H, W), dtype=torch.float64, device=t.device)
    with torch.cuda.device(t.device):
        _lib.check(lib.uwie_prim_transmission(_p(t), _DT[t.dtype], H, W, _p(light), float(omega), int(r), float(eps), float(guard),
                                              int(bool(clip_t0)), _p(out), _p(ws), ws.numel(), _stream(t)))
    return out


def recover(img, t_plane, light, out_dtype=torch.float64):
    """clip((img - A) / t + A, 0, 1) stored as out_dtype (ES:237-249 float64; SS:182-188 float32)."""
    t = _float_image(img)
    H, W, _ = t.shape
    tp = as_device(t_plane).to(torch.float64).contiguous()
    light = as_device(light).to(t.dtype).reshape(-1)[:3].contiguous()
    out = torch.empty((H, W, 3), dtype=out_dtype, device=t.device)
    with torch.cuda.device(t.device):
        _lib.check(_lib.load().uwie_prim_recover(_p(t), _DT[t.dtype], H, W, _p(tp), _p(light), _p(out), _DT[out_dtype], _stream(t)))
    return out


def percentile_stretch(img, lo, hi, guard):
    """Per-channel clip((x - p_lo) / (p_hi - p_lo + guard), 0, 1) with exact np.percentile (ES:252-273, SS:190-219)."""
    t = _float_image(img)
    H, W, _ = t.shape
    lib = _lib.load()
    ws = _ws(lib.uwie_prim_percentile_stretch_workspace_bytes(_DT[t.dtype]), t.device)
    out = torch.empty_like(t)
    with torch.cuda.device(t.device):
        _lib.check(lib.uwie_prim_percentile_stretch(_p(t), _DT[t.dtype], H, W, float(lo), float(hi), float(guard), _p(out),
                                                    _p(ws), ws.numel(), _stream(t)))
    return out


def gamma(img, exponent, clip):
    t = as_device(img)
    if t.dtype not in _DT:
        raise ValueError("gamma takes float32 / float64")
    out = torch.empty_like(t)
    with torch.cuda.device(t.device):
        _lib.check(_lib.load().uwie_prim_gamma(_p(t), _DT[t.dtype], t.numel(), float(exponent), int(bool(clip)), _p(out), _stream(t)))
    return out


def scale_clip(img, scale3):
    t = _float_image(img)
    H, W, _ = t.shape
    arr = (C.c_double * 3)(*[float(v) for v in scale3])
    out = torch.empty_like(t)
    with torch.cuda.device(t.device):
        _lib.check(_lib.load().uwie_prim_scale_clip(_p(t), _DT[t.dtype], H, W, arr, _p(out), _stream(t)))
    return out


def channel_means(img):
    t = _float_image(img)
    H, W, _ = t.shape
    ws = _ws(4096, t.device)
    out = torch.empty(3, dtype=torch.float64, device=t.device)
    with torch.cuda.device(t.device):
        _lib.check(_lib.load().uwie_prim_channel_means(_p(t), _DT[t.dtype], H, W, _p(out), _p(ws), ws.numel(), _stream(t)))
    return out


def region_score(block) -> float:
    """compute_Q(block) (ES:147-188, SS:115-157) -> Python float."""
    t = _float_image(block)
    H, W, _ = t.shape
    lib = _lib.load()
    ws = _ws(lib.uwie_prim_region_score_workspace_bytes(H, W), t.device)
    out = torch.empty(1, dtype=torch.float64, device=t.device)
    with torch.cuda.device(t.device):
        _lib.check(lib.uwie_prim_region_score(_p(t), _DT[t.dtype], H, W, _p(out), _p(ws), ws.numel(), _stream(t)))
    return float(out.item())


def brightest_pixel(block):
    """get_brightest_pixel(block) (ES:191-206, SS:159-165) -> (3,) tensor of the block's dtype."""
    t = _float_image(block)
    H, W, _ = t.shape
    ws = _ws(256, t.device)
    out = torch.empty(3, dtype=t.dtype, device=t.device)
    with torch.cuda.device(t.device):
        _lib.check(_lib.load().uwie_prim_brightest_pixel(_p(t), _DT[t.dtype], H, W, _p(out), _p(ws), ws.numel(), _stream(t)))
    return out
