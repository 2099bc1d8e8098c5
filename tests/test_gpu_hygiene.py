"""Memory hygiene without compute-sanitizer (the tool is closed on this GPU pool: "compute-sanitizer is closed on this
pool and stays closed"; gpurun refuses it).  Two in-process substitutes, wired into every scratch allocation of the
Python faces (_lib.scratch):

  * initcheck's question — does any result depend on scratch memory that was never written?  Every workspace, the
    engine's included (its Canny rim-label / node arrays are deliberately never cleared), is filled with 0x00, 0xFF and 0xA5 before
    use; all results must be bit-identical across the three fills.
  * memcheck's question — does any kernel write outside the bytes `*_workspace_bytes` promised?  Every workspace sits
    between two 4 KiB canary bands that must come back untouched.

racecheck's question is covered by determinism: the same inputs give the same bits run after run, on one lane or through
graph replays (tests/test_gpu_parity.py::test_batch_slots_are_independent_and_deterministic, ::test_graph_replay_*).
"""
from __future__ import annotations

import os

import cv2
import numpy as np
import pytest
import torch

from underwater_image_enhancement_b200 import _lib, features, prims, quality, vgg
from underwater_image_enhancement_b200.engine import Engine
from underwater_image_enhancement_b200.synth import synth_image

pytestmark = pytest.mark.gpu


def _everything():
    """One small call of every entry point; returns a flat dict of NumPy results."""
    out = {}
    h, w = 100, 150
    frames = np.stack([synth_image(s, h, w) for s in (0, 1, 2)])
    d = torch.from_numpy(frames).cuda()
    for lib in ("six", "es"):
        eng = Engine(lib, h, w, max_batch=2)
        r = eng.run(d[:2], want_all=True, want_best=True)
        eng.check()
        for k, v in r.items():
            out[f"{lib}.dev.{k}"] = v.cpu().numpy()
        rh = eng.run_host(frames, want_all=True, want_best=True)
        for k, v in rh.items():
            out[f"{lib}.host.{k}"] = np.array(v)
        eng.close()
    out["qa"] = quality.quality_scores_batch(d).cpu().numpy()
    out["fe"] = features.extract_all_features_batch(d).cpu().numpy()
    img = torch.from_numpy(np.ascontiguousarray((frames.astype(np.float32) / 255.0).transpose(0, 3, 1, 2))).cuda()
    p = torch.tensor([[5.0, 10.0, 2.0], [95.0, 90.0, 98.0], [0.5, 0.7, 0.4], [1.2, 0.9, 1.4]], device="cuda")
    out["vgg.f32"] = vgg.DifferentiableEnhancement().forward_dev(img, p, True)[0].cpu().numpy()
    out["vgg.u8"] = vgg.enhance_frames_u8(d, p, True)[0].cpu().numpy()
    out["vgg.feat"] = vgg.extract_all_features(d).cpu().numpy()
    out["vgg.pre"] = vgg.preprocess_for_vgg(d, 64).cpu().numpy()
    planes = np.stack([cv2.cvtColor(f, cv2.COLOR_RGB2LAB)[:, :, 0] for f in frames])
    out["clahe"] = prims.clahe_lab(planes, 2.0, (8, 8)).cpu().numpy()
    out["clahe.split"] = prims.clahe_lab(planes[:1], 3.0, (2, 2)).cpu().numpy()     # few tiles -> slabs + last-block LUT
    out["eq"] = prims.equalize_hist(frames).cpu().numpy()
    light = torch.tensor([[0.3, 0.6, 0.7]] * 3, device="cuda")
    out["thead"] = prims.transmission_head_u8(frames, light, 0.5, 1e-6, True).cpu().numpy()
    rng = np.random.default_rng(0)
    for dt in (np.float32, np.float64):
        x = rng.random((h, w, 3)).astype(dt)
        A = prims.atmospheric_light(x)
        t = prims.transmission(x, A, 0.5, 15, 1e-3, 1e-10, False)
        rec = prims.recover(x, t, A, torch.float64 if dt == np.float64 else torch.float32)
        st = prims.percentile_stretch(rec, 10, 95, 1e-10)
        tag = dt.__name__
        out[f"{tag}.A"], out[f"{tag}.t"], out[f"{tag}.st"] = A.cpu().numpy(), t.cpu().numpy(), st.cpu().numpy()
        out[f"{tag}.q"] = np.array([prims.region_score(x)])
        out[f"{tag}.means"] = prims.channel_means(x).cpu().numpy()
    return out


def test_results_do_not_depend_on_scratch_contents_and_stay_inside_their_workspaces(cuda_device):
    saved = {k: os.environ.get(k) for k in ("UWIE_POISON", "UWIE_GUARD")}
    runs = {}
    try:
        os.environ["UWIE_GUARD"] = "4096"
        for fill in (0x00, 0xFF, 0xA5):
            os.environ["UWIE_POISON"] = str(fill)
            runs[fill] = _everything()
            assert _lib.guards_intact(), f"a kernel wrote outside its workspace (scratch fill {fill:#x})"
    finally:
        for k, v in saved.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    base = runs[0x00]
    for fill in (0xFF, 0xA5):
        for k, v in base.items():
            if not np.array_equal(v, runs[fill][k], equal_nan=True):
                w = runs[fill][k]
                where = np.argwhere(~((v == w) | (np.isnan(v) & np.isnan(w))))[:8].tolist() if v.shape == w.shape else "shapes"
                raise AssertionError(f"{k} depends on the initial scratch contents ({fill:#x} vs 0x00) at {where}")


def test_caller_owned_outputs_are_not_overrun(cuda_device):
    """Outputs placed between canary bands: the engine writes exactly the bytes the header documents."""
    h, w, B, S = 67, 121, 2, 6                                  # odd sizes: no kernel may round a store up past the end
    frames = torch.from_numpy(np.stack([synth_image(s, h, w) for s in (3, 4)])).cuda()
    G = 1024

    def banded(shape, dtype):
        n = int(np.prod(shape)) * torch.empty((), dtype=dtype).element_size()
        raw = torch.full((n + 2 * G,), 0xC3, dtype=torch.uint8, device="cuda")
        return raw, raw[G:G + n].view(dtype).view(shape)

    bufs = {"out_all": banded((B, S, h, w, 3), torch.uint8), "out_best": banded((B, h, w, 3), torch.uint8),
            "scores": banded((B, S, 9), torch.float64), "best": banded((B,), torch.int32),
            "types": banded((B,), torch.int32), "light": banded((B, 3), torch.float32)}
    eng = Engine("six", h, w, max_batch=B)
    eng.run(frames, want_all=True, want_best=True, **{k: v[1] for k, v in bufs.items()})
    eng.check()
    eng.close()
    for k, (raw, view) in bufs.items():
        assert bool((raw[:G] == 0xC3).all()) and bool((raw[-G:] == 0xC3).all()), k
        assert view.numel() > 0
