"""Training-time face of the enhancer: DifferentiableEnhancement as a torch.autograd.Function
whose forward AND backward are the library's CUDA kernels (uwie_vgg_enhance_dev / uwie_vgg_enhance_bwd).

Gradients follow what autograd derives for vgg_16_UIE.py:32-128: omega and gamma receive gradients; L_low and
L_high do not (the reference turns them into integer ranks through .item(), VGG:78-79, which cuts the graph).
The image's own gradient is not built (the training loops feed data, vgg_16_UIE.py:529): asking for it raises.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib


def _col(x, batch, dev, name):
    t = x if isinstance(x, torch.Tensor) else torch.as_tensor(x, dtype=torch.float32)
    t = t.to(dev, torch.float32).reshape(-1)
    if t.numel() == 1 and batch > 1:
        t = t.expand(batch)
    if t.numel() != batch:
        raise ValueError(f"{name}: expected {batch} values, got {t.numel()}")
    return t


class EnhanceFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, img, l_low, l_high, omega, gamma):
        lib = _lib.load()
        img = img.contiguous().to(torch.float32)
        B, _, H, W = img.shape
        params = torch.stack([l_low.detach(), l_high.detach(), omega.detach(), gamma.detach()]).contiguous()
        need = lib.uwie_vgg_workspace_bytes(B)
        ws = _lib.scratch(need, img.device)
        out = torch.empty_like(img)
        stream = C.c_void_p(torch.cuda.current_stream(img.device).cuda_stream)
        with torch.cuda.device(img.device):
            _lib.check(lib.uwie_vgg_enhance_dev(C.c_void_p(img.data_ptr()), B, H, W, C.c_void_p(params.data_ptr()), 0, None,
                                                C.c_void_p(out.data_ptr()), C.c_void_p(ws.data_ptr()), need, stream))
        ctx.save_for_backward(img, ws)
        ctx.shapes = (omega.shape, gamma.shape)
        return out

    @staticmethod
    def backward(ctx, grad_out):
        if ctx.needs_input_grad[0]:
            raise NotImplementedError("DifferentiableEnhancement: the gradient with respect to the image is not built "
                                      "(the reference's training loops feed data there)")
        img, ws = ctx.saved_tensors
        B, _, H, W = img.shape
        lib = _lib.load()
        g = torch.empty((2, B), dtype=torch.float64, device=img.device)
        go = grad_out.contiguous().to(torch.float32)
        stream = C.c_void_p(torch.cuda.current_stream(img.device).cuda_stream)
        with torch.cuda.device(img.device):
            _lib.check(lib.uwie_vgg_enhance_bwd(C.c_void_p(img.data_ptr()), C.c_void_p(go.data_ptr()), B, H, W,
                                                C.c_void_p(ws.data_ptr()), ws.numel(), C.c_void_p(g.data_ptr()), stream))
        g = g.to(torch.float32)
        return None, None, None, g[0].reshape(ctx.shapes[0]), g[1].reshape(ctx.shapes[1])


def enhance_with_grad(img: torch.Tensor, params: dict) -> torch.Tensor:
    """DifferentiableEnhancement.forward with a graph (vgg_16_UIE.py:32-55; 'omega' and 'gamma' must be present: the
    training loops always pass all four parameters, vgg_16_UIE.py:520-529)."""
    if img.dim() != 4 or img.shape[1] != 3:
        raise ValueError(f"img must be (B, 3, H, W), got {tuple(img.shape)}")
    if "omega" not in params or "gamma" not in params:
        raise ValueError("the differentiable path needs all four parameters (L_low, L_high, omega, gamma)")
    B, dev = img.shape[0], img.device
    lo, hi = _col(params["L_low"], B, dev, "L_low"), _col(params["L_high"], B, dev, "L_high")
    om, ga = params["omega"], params["gamma"]
    om_flat = om.to(dev, torch.float32).reshape(-1) if isinstance(om, torch.Tensor) else _col(om, B, dev, "omega")
    ga_flat = ga.to(dev, torch.float32).reshape(-1) if isinstance(ga, torch.Tensor) else _col(ga, B, dev, "gamma")
    if om_flat.numel() != B or ga_flat.numel() != B:
        raise ValueError("omega / gamma: one value per image")
    return EnhanceFunction.apply(img, lo, hi, om_flat, ga_flat)
