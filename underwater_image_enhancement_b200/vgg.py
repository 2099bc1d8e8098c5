"""Host-side mirror of the VGG-guided path's per-pixel pieces (use_trained_model.py / vgg_16_UIE.py).

Same names and argument meaning as the reference; all arithmetic runs in libuwie_b200.so (uwie_vgg_*),
PyTorch only provides device memory and the current stream.  The network (ImprovedVGGParameterNet)
is not part of this package: it stays stock PyTorch (out of scope).
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib


def _stream(dev):
    return C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


def _workspace(batch, dev):
    need = _lib.load().uwie_vgg_workspace_bytes(int(batch))
    return _lib.scratch(need, dev), need


def _floats(x, batch, name):
    """(B,1) / (B,) tensor, sequence or scalar -> ctypes float array of length B (host side, like .item())."""
    if x is None:
        return None
    if isinstance(x, torch.Tensor):
        x = x.detach().reshape(-1).to("cpu", torch.float32).numpy()
    arr = np.asarray(x, dtype=np.float32).reshape(-1)
    if arr.size == 1 and batch > 1:
        arr = np.repeat(arr, batch)
    if arr.size != batch:
        raise ValueError(f"{name}: expected {batch} values, got {arr.size}")
    return (C.c_float * batch)(*arr.tolist())


def _needs_grad(img, params) -> bool:
    if not torch.is_grad_enabled():
        return False
    ts = [img] + [v for v in params.values() if isinstance(v, torch.Tensor)]
    return any(t.requires_grad for t in ts)


class DifferentiableEnhancement:
    """vgg_16_UIE.py:24-128: forward(img, params) on a CUDA float32 NCHW batch.

    Inference (no gradient wanted) runs the fused selection + enhancement kernels.  When autograd is recording and the
    image or a parameter requires grad (the training loops, vgg_16_UIE.py:529 / deep_learning_parameters.py:281), the
    call goes through autograd_enhance.EnhanceFunction, whose backward is a CUDA kernel too; it never silently returns
    a graph-less tensor."""

    def to(self, *a, **k):            # nn.Module look-alikes used by use_trained_model.py:27
        return self

    def eval(self):
        return self

    def train(self, mode=True):
        return self

    def forward(self, img: torch.Tensor, params: dict) -> torch.Tensor:
        if not (isinstance(img, torch.Tensor) and img.is_cuda):
            raise ValueError("img must be a CUDA tensor: the enhancement has no CPU path")
        if _needs_grad(img, params):
            from .autograd_enhance import enhance_with_grad
            return enhance_with_grad(img, params)
        if img.dim() != 4 or img.shape[1] != 3:
            raise ValueError(f"img must be (B, 3, H, W), got {tuple(img.shape)}")
        img = img.contiguous().to(torch.float32)
        B, _, H, W = img.shape
        lo = _floats(params["L_low"], B, "L_low")                      # KeyError like the reference
        hi = _floats(params["L_high"], B, "L_high")
        om = _floats(params.get("omega"), B, "omega")
        ga = _floats(params.get("gamma"), B, "gamma")
        out = torch.empty_like(img)
        ws, need = _workspace(B, img.device)
        with torch.cuda.device(img.device):
            _lib.check(_lib.load().uwie_vgg_enhance(C.c_void_p(img.data_ptr()), B, H, W, lo, hi, om, ga,
                                                    C.c_void_p(out.data_ptr()), C.c_void_p(ws.data_ptr()), need,
                                                    _stream(img.device)))
        ws.record_stream(torch.cuda.current_stream(img.device))
        return out

    __call__ = forward

    def forward_dev(self, img: torch.Tensor, params_dev: torch.Tensor, clamp_ranges: bool = False):
        """Same enhancement with the parameters resident on the device: params_dev float32 [4, B] = rows L_low, L_high,
        omega, gamma (uwie_vgg_enhance_dev).  Returns (enhanced, used [4, B]); nothing synchronises."""
        if not (img.is_cuda and params_dev.is_cuda):
            raise ValueError("img and params_dev must be CUDA tensors")
        img = img.contiguous().to(torch.float32)
        B, _, H, W = img.shape
        params_dev = params_dev.to(torch.float32).contiguous()
        if tuple(params_dev.shape) != (4, B):
            raise ValueError(f"params_dev must be [4, {B}], got {tuple(params_dev.shape)}")
        out = torch.empty_like(img)
        used = torch.empty_like(params_dev)
        ws, need = _workspace(B, img.device)
        with torch.cuda.device(img.device):
            _lib.check(_lib.load().uwie_vgg_enhance_dev(C.c_void_p(img.data_ptr()), B, H, W, C.c_void_p(params_dev.data_ptr()),
                                                        int(bool(clamp_ranges)), C.c_void_p(used.data_ptr()),
                                                        C.c_void_p(out.data_ptr()), C.c_void_p(ws.data_ptr()), need,
                                                        _stream(img.device)))
        ws.record_stream(torch.cuda.current_stream(img.device))
        return out, used


def enhance_frames_u8(frames: torch.Tensor, params_dev: torch.Tensor, clamp_ranges: bool = False):
    """uint8 CUDA frames [B, H, W, 3] -> (float32 [B, 3, H, W] enhanced, float32 [4, B] parameters applied): the
    enhancement of frame / 255 with device-resident parameters, reading the frame as bytes (uwie_vgg_enhance_u8)."""
    if not (frames.is_cuda and params_dev.is_cuda and frames.dtype == torch.uint8):
        raise ValueError("frames must be a uint8 CUDA tensor and params_dev a CUDA tensor")
    frames = frames.contiguous()
    B, H, W, _ = frames.shape
    params_dev = params_dev.to(torch.float32).contiguous()
    if tuple(params_dev.shape) != (4, B):
        raise ValueError(f"params_dev must be [4, {B}], got {tuple(params_dev.shape)}")
    out = torch.empty((B, 3, H, W), dtype=torch.float32, device=frames.device)
    used = torch.empty_like(params_dev)
    ws, need = _workspace(B, frames.device)
    with torch.cuda.device(frames.device):
        _lib.check(_lib.load().uwie_vgg_enhance_u8(C.c_void_p(frames.data_ptr()), B, H, W, C.c_void_p(params_dev.data_ptr()),
                                                   int(bool(clamp_ranges)), C.c_void_p(used.data_ptr()),
                                                   C.c_void_p(out.data_ptr()), C.c_void_p(ws.data_ptr()), need,
                                                   _stream(frames.device)))
    ws.record_stream(torch.cuda.current_stream(frames.device))
    return out, used


def _frames_u8(img, device):
    """HWC or BHWC uint8 image(s), NumPy or torch -> contiguous CUDA uint8 [B,H,W,3]."""
    t = torch.as_tensor(img) if not isinstance(img, torch.Tensor) else img
    if t.dim() == 3:
        t = t[None]
    if t.dim() != 4 or t.shape[-1] != 3:
        raise ValueError(f"expected (H, W, 3) or (B, H, W, 3), got {tuple(t.shape)}")
    return t.to(device).contiguous()


def extract_all_features(img, device="cuda:0") -> torch.Tensor:
    """vgg_16_UIE.py:435-466 for uint8 RGB frame(s): float32 [B, 79] on the device.

    The reference's only caller (use_trained_model.py:55) passes imread / 255; pass the bytes themselves
    (the reference divides uint8 input by 255 itself, VGG:425-433)."""
    t = torch.as_tensor(img) if not isinstance(img, torch.Tensor) else img
    if t.dtype != torch.uint8:
        raise TypeError("extract_all_features takes the uint8 frame (float32 k/255 images: pass the bytes)")
    frames = _frames_u8(t, device)
    B, H, W, _ = frames.shape
    feat = torch.empty((B, _lib.VGG_FEATURES), dtype=torch.float32, device=frames.device)
    ws, need = _workspace(B, frames.device)
    with torch.cuda.device(frames.device):
        _lib.check(_lib.load().uwie_vgg_features_u8(C.c_void_p(frames.data_ptr()), B, H, W, C.c_void_p(feat.data_ptr()),
                                                    C.c_void_p(ws.data_ptr()), need, _stream(frames.device)))
    ws.record_stream(torch.cuda.current_stream(frames.device))
    return feat


def preprocess_for_vgg(img, input_size: int = 224, device="cuda:0") -> torch.Tensor:
    """use_trained_model.py:39-46 for uint8 or float32 [0,1] HWC frame(s): float32 [B, 3, size, size]."""
    t = torch.as_tensor(img) if not isinstance(img, torch.Tensor) else img
    if t.dim() == 3:
        t = t[None]
    if t.dim() != 4 or t.shape[-1] != 3:
        raise ValueError(f"expected (H, W, 3) or (B, H, W, 3), got {tuple(t.shape)}")
    if t.dtype == torch.uint8:
        fn = _lib.load().uwie_vgg_preprocess_u8
    elif t.dtype == torch.float32:
        fn = _lib.load().uwie_vgg_preprocess_f32
    else:
        raise TypeError(f"unsupported dtype {t.dtype}")
    t = t.to(device).contiguous()
    B, H, W, _ = t.shape
    out = torch.empty((B, 3, input_size, input_size), dtype=torch.float32, device=t.device)
    with torch.cuda.device(t.device):
        _lib.check(fn(C.c_void_p(t.data_ptr()), B, H, W, int(input_size), C.c_void_p(out.data_ptr()), _stream(t.device)))
    return out
