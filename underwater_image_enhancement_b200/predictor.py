"""EnhancementPredictor of use_trained_model.py:13-164, batched (BASELINE.json configs[3]).

The parameter network (ImprovedVGGParameterNet, vgg_16_UIE.py:135-250) is the reference's own class and stays stock
PyTorch (north_star: out of scope) — it is imported from the maintainer's tree, never restated here.  Everything
around it runs in libuwie_b200.so on whole batches and without host round trips:

    uint8 frames [B,H,W,3] on the device
      -> uwie_vgg_features_u8   18 statistics zero-padded to 79            (vgg_16_UIE.py:435-466)
      -> uwie_vgg_preprocess_u8 224x224 fixed-point resize + normalisation (use_trained_model.py:39-46)
      -> network (PyTorch)      {omega, gamma, L_low, L_high} as [B,1] tensors, left on the device
      -> uwie_vgg_enhance_u8    safety clamps of UTM:74-79, order statistics from the byte histograms, stretch, dehaze,
                                gamma on the device; float32 NCHW out

The reference's per-image API (predict_parameters -> dict of floats, enhance_image -> float32 HWC array,
process_single_image, process_folder) is kept on top of the batched calls.
"""
from __future__ import annotations

import importlib
import os
import sys
from pathlib import Path

import numpy as np
import torch

from . import vgg

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PARAM_ORDER = ("L_low", "L_high", "omega", "gamma")          # rows of the device parameter block


def load_network_class():
    """ImprovedVGGParameterNet from the reference's vgg_16_UIE.py: whatever `import vgg_16_UIE` finds on sys.path (the
    maintainer's tree), else UWIE_REF_DIR, baseline/_ref, /root/reference."""
    try:
        mod = importlib.import_module("vgg_16_UIE")
    except ImportError:
        for d in (os.environ.get("UWIE_REF_DIR", ""), os.path.join(_ROOT, "baseline", "_ref"), "/root/reference"):
            if d and os.path.exists(os.path.join(d, "vgg_16_UIE.py")):
                sys.path.append(d)
                mod = importlib.import_module("vgg_16_UIE")
                break
        else:
            raise
    return mod.ImprovedVGGParameterNet


class EnhancementPredictor:
    def __init__(self, model_path=None, device="cuda", input_size=224, model=None, net_dtype=torch.float32):
        if not torch.cuda.is_available():
            raise RuntimeError("uwie: no CUDA device — the VGG-guided path has no CPU fallback")
        self.device = torch.device("cuda:0" if device in ("cuda", "cpu") else device)   # UTM:15: 'cuda' when available
        self.input_size = int(input_size)
        self.net_dtype = net_dtype
        if model is None:
            print(f"載入模型: {model_path}  (device={self.device})")
            model = load_network_class()(pretrained=False, hidden_dim=256, use_features=True)          # UTM:21
            ckpt = torch.load(model_path, map_location="cpu")
            model.load_state_dict(ckpt.get("model_state_dict", ckpt))                                  # UTM:23-25
        self.model = model.to(self.device).eval()
        self.enhancer = vgg.DifferentiableEnhancement()
        print("✓ 模型與增強模組已載入")

    # ------------------------------------------------------------------ batched device path
    def _frames(self, img):
        a = img if isinstance(img, torch.Tensor) else torch.as_tensor(np.ascontiguousarray(img))
        if a.dim() == 3:
            a = a[None]
        if a.dtype != torch.uint8:                               # the reference's float32 k / 255 images -> their bytes
            a = (a.to(self.device, torch.float32) * 255).to(torch.uint8)      # (img * 255).astype(uint8), UTM:41
        return a.to(self.device).contiguous()

    @torch.no_grad()
    def predict_parameters_batch(self, frames: torch.Tensor) -> torch.Tensor:
        """uint8 [B,H,W,3] on the device -> float32 [4, B] (rows L_low, L_high, omega, gamma), unclamped, on the device."""
        feats = vgg.extract_all_features(frames, device=self.device)
        x = vgg.preprocess_for_vgg(frames, self.input_size, device=self.device)
        if self.net_dtype != torch.float32:
            with torch.autocast("cuda", dtype=self.net_dtype):
                out = self.model(x, feats)
        else:
            out = self.model(x, feats)
        return torch.stack([out[k].reshape(-1).float() for k in PARAM_ORDER]).contiguous()

    @torch.no_grad()
    def enhance_batch(self, frames: torch.Tensor, params_dev: torch.Tensor | None = None):
        """uint8 [B,H,W,3] -> (float32 [B,3,H,W] enhanced, float32 [4,B] parameters applied), all on the device."""
        frames = self._frames(frames)
        if params_dev is None:
            params_dev = self.predict_parameters_batch(frames)
        return vgg.enhance_frames_u8(frames, params_dev, clamp_ranges=True)              # frame / 255 (UTM:119) inside

    # ------------------------------------------------------------------ the reference's per-image API
    def predict_parameters(self, img):
        """UTM:53-81: dict of Python floats, clamped, with the extra keys the reference adds."""
        frames = self._frames(img)
        p = self.predict_parameters_batch(frames)[:, 0].cpu().numpy().astype(np.float64)
        params = {k: float(p[i]) for i, k in enumerate(PARAM_ORDER)}
        params.setdefault("guided_radius", 15.0)
        params.setdefault("use_gamma", 1.0)
        params["omega"] = float(np.clip(params.get("omega", 0.6), 0.1, 0.9))
        params["gamma"] = float(np.clip(params.get("gamma", 1.2), 0.5, 3.0))
        params["L_low"] = float(np.clip(params.get("L_low", 10.0), 1.0, 30.0))
        params["L_high"] = float(np.clip(params.get("L_high", 90.0), 65.0, 99.0))
        params["guided_radius"] = float(np.clip(params["guided_radius"], 1.0, 50.0))
        params["use_gamma"] = float(np.clip(params["use_gamma"], 0.0, 1.0))
        return params

    def enhance_image(self, img, params=None):
        """UTM:83-111: float32 (H, W, 3) in [0, 1]."""
        if params is None:
            params = self.predict_parameters(img)
        a = np.asarray(img)
        if a.dtype == np.uint8:
            a = a.astype(np.float32) / np.float32(255.0)
        t = torch.from_numpy(a.astype(np.float32)).permute(2, 0, 1)[None].to(self.device).contiguous()
        p = {k: torch.tensor([[float(params[k])]], dtype=torch.float32) for k in ("omega", "gamma", "L_low", "L_high")}
        with torch.no_grad():
            out = self.enhancer(t, p)
        enhanced = out[0].permute(1, 2, 0).cpu().numpy()
        enhanced = np.clip(enhanced, 0.0, 1.0)
        if not np.isfinite(enhanced).all():
            enhanced = np.clip(np.nan_to_num(enhanced, nan=0.0, posinf=1.0, neginf=0.0), 0.0, 1.0)
        return enhanced

    def process_single_image(self, input_path, output_path=None, show_params=True):
        """UTM:113-146."""
        import cv2
        input_path = Path(input_path)
        img = cv2.imread(str(input_path))
        if img is None:
            raise ValueError(f"無法讀取影像: {input_path}")
        img = cv2.cvtColor(img, cv2.COLOR_BGR2RGB).astype(np.float32) / 255.0
        params = self.predict_parameters(img)
        if show_params:
            print("\n預測的最佳參數:")
            for k, v in params.items():
                print(f"  {k}: {v:.4f}")
        enhanced = self.enhance_image(img, params)
        enhanced_uint8 = (np.clip(enhanced, 0, 1) * 255).astype(np.uint8)
        if output_path is None:
            output_path = input_path.parent / f"{input_path.stem}_enhanced.png"
        else:
            output_path = Path(output_path)
            if output_path.suffix == "":
                output_path.mkdir(parents=True, exist_ok=True)
                output_path = output_path / f"{input_path.stem}_enhanced.png"
            else:
                output_path.parent.mkdir(parents=True, exist_ok=True)
        cv2.imwrite(str(output_path), cv2.cvtColor(enhanced_uint8, cv2.COLOR_RGB2BGR))
        print(f"儲存: {output_path}")
        return enhanced, params

    def process_folder(self, input_folder, output_folder, batch: int = 16, workers: int = 4):
        """UTM:148-164, streamed: files of one size go through the batched device path `batch` at a time while the
        decode of the next and the encode of the previous batch run on worker threads (frame_io.py)."""
        from .frame_io import FrameReader, FrameWriter, list_images
        out_dir = Path(output_folder)
        out_dir.mkdir(parents=True, exist_ok=True)
        files = list_images(input_folder, (".png", ".jpg", ".jpeg"))
        if not files:
            print("找不到任何影像檔案！")
            return
        reader, writer = FrameReader(files, batch=batch, workers=workers), FrameWriter(workers=workers)
        done = 0
        try:
            for paths, frames in reader:
                n, h, w = frames.shape[:3]
                try:
                    d = torch.from_numpy(frames).to(self.device, non_blocking=True)
                    enhanced, _ = self.enhance_batch(d)
                    q = (enhanced.clamp(0, 1) * 255).to(torch.uint8).permute(0, 2, 3, 1).contiguous()   # UTM:126
                    slot = writer.slot((batch, h, w, 3))
                    torch.from_numpy(slot)[:n].copy_(q)
                except Exception as e:                          # noqa: BLE001  (UTM:163-164)
                    for p in paths:
                        print(f"  失敗: {p.name}: {e}")
                    continue
                for i, p in enumerate(paths):
                    done += 1
                    print(f"[{done}/{len(files)}] 處理: {p.name}")
                    writer.write(out_dir / f"{p.stem}_enhanced.png", slot[i])
            for p in reader.failed:
                print(f"  失敗: 無法讀取影像: {p}")
        finally:
            writer.close()
            reader.close()
