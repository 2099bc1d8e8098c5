"""The drop-in boundary with the reference's own names and the "next" rows built on it:

  * the reference's example_usage.py (baseline/_ref, unmodified) runs against the shim modules;
  * non byte-valued float images (np.random.rand, example_usage.py:27/44/112) through the primitives vs the oracle;
  * six_stadigy's file loop with the pinned double-buffered reader / writer (f-1) and its CSV schema;
  * SelfSupervisedSystem.build_dataset fed from the device (f-2);
  * EnhancementPredictor batched vs the reference's own predictor on a random-init checkpoint (f-3);
  * the autograd-capable enhancer vs torch autograd on the reference's DifferentiableEnhancement (f-4).
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
import types

import numpy as np
import pytest
import torch

from oracle import restatement as o
from underwater_image_enhancement_b200 import compat
from underwater_image_enhancement_b200.synth import synth_image

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref")
SHIMS = os.path.join(ROOT, "shims")
needs_ref = pytest.mark.skipif(not os.path.exists(os.path.join(REF, "example_usage.py")),
                               reason="baseline/_ref (verbatim copy of the reference, made by __graft_entry__.build()) is absent")


def _deploy(tmp_path):
    """What a maintainer does: the reference tree with its four hot-path modules replaced by the shims."""
    d = tmp_path / "tree"
    d.mkdir()
    for name in ("example_usage.py", "config.py", "main.py", "vgg_16_UIE.py"):
        shutil.copyfile(os.path.join(REF, name), d / name)
    for name in os.listdir(SHIMS):
        if name.endswith(".py"):
            shutil.copyfile(os.path.join(SHIMS, name), d / name)
    return d


@needs_ref
def test_reference_example_usage_runs_unchanged(cuda_device, tmp_path):
    d = _deploy(tmp_path)
    env = dict(os.environ, PYTHONPATH=ROOT, PYTHONIOENCODING="utf-8")
    r = subprocess.run([sys.executable, "example_usage.py", "--all"], cwd=d, env=env, capture_output=True, text=True, timeout=600)
    out = r.stdout + r.stderr
    assert r.returncode == 0, out[-3000:]
    assert "執行失敗" not in out and "範例執行失敗" not in out and "Traceback" not in out, out[-3000:]
    assert "✓ 策略 StrongDehazing 執行成功" in out                      # example 1: np.random.rand image through apply_strategy
    for name in ("StrongDehazing", "MediumDehazing", "LightEnhancement", "CLAHEEnhancement", "HistogramEqualization"):
        assert f"✓ {name}" in out, name                                 # example 2
    assert "特徵維度: 79" in out and "✓ 特徵品質良好" in out            # example 4
    assert "最佳策略:" in out                                            # example 5
    assert "所有範例執行完成" in out


@needs_ref
def test_shim_module_self_tests(cuda_device, tmp_path):
    d = _deploy(tmp_path)
    env = dict(os.environ, PYTHONPATH=ROOT, PYTHONIOENCODING="utf-8", MPLBACKEND="Agg")
    code = ("import enhancement_strategies as e, quality_assessment as q, feature_extraction as f\n"
            "r = e.test_strategies(); assert len(r) == 5 and all(v.shape == (256, 256, 3) for v in r.values())\n"
            "s = q.test_quality_assessment(); assert s['Low Contrast'][1]['contrast'] == 0.0 and s['Random'][1]['sharpness'] == 50.0\n"
            "v = f.test_feature_extraction(); assert len(v) == 79\n")
    r = subprocess.run([sys.executable, "-c", code], cwd=d, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout + r.stderr)[-3000:]


# ------------------------------------------------------------------ float images that are not made from bytes
@pytest.mark.parametrize("dt", [np.float64, np.float32])
def test_es_apply_strategy_on_random_float_images(cuda_device, dt):
    """example_usage.py:27, 112: np.random.rand(256, 256, 3) and its 0.7 x + 0.15 version through every strategy of
    config.py.  Float stages: within 1/255 of the reference's float image (north_star); the strategies that quantise
    first (CLAHE, equalizeHist) agree to float64 rounding."""
    rng = np.random.default_rng(12)
    for img in (rng.random((256, 256, 3)).astype(dt), (rng.random((256, 256, 3)) * 0.7 + 0.15).astype(dt)):
        assert compat.as_frame(img) is None
        for name, params in o.ES_DEFAULT_STRATEGIES.items():
            got = compat.ESApi.apply_strategy(img, name, params)
            want = o.es_strategy(img, name, params)
            assert got.dtype == np.float64 and got.shape == img.shape, name
            d = np.abs(got - want)
            assert d.max() <= 1.0 / 255.0, f"{name}: max |d| = {d.max():.3e}"
            if name in ("clahe_enhancement", "histogram_equalization"):
                assert d.max() < 1e-12, name


def test_es_static_methods_match_reference_types(cuda_device):
    rng = np.random.default_rng(3)
    img = rng.random((96, 128, 3))
    ES = compat.ESApi
    A = ES.estimate_atmospheric_light(img)
    assert A.shape == img.shape and A.dtype == np.float64 and np.all(A == A[0, 0])
    assert np.array_equal(A[0, 0], o.quadtree_airlight(img, 1))
    t = ES.estimate_transmission(img, A, omega=0.5, r=15)
    assert t.shape == img.shape[:2] and t.dtype == np.float64 and t.min() >= 0.1 and t.max() <= 1.0
    rec = ES.recover_image(img, t, A)
    assert np.abs(rec - o.es_recover(img, t, A)).max() < 1e-12
    assert abs(ES.compute_Q(img) - o.region_score(img)) < 1e-9
    assert np.array_equal(ES.get_brightest_pixel(img), img.reshape(-1, 3)[np.argmax(img.sum(axis=2))])
    assert np.array_equal(ES.clahe_enhancement(img), o.clahe_lab_u8(img, 2.0, (8, 8)).astype(np.float64) / 255.0)
    assert np.array_equal(ES.histogram_equalization(img), o.es_equalize(img))
    assert np.array_equal(ES.color_enhancement(img, 15, 95), o.percentile_stretch(img, 15, 95, 1e-10))
    assert np.abs(ES.gamma_correction(img, 1.2) - o.es_gamma(img, 1.2)).max() < 1e-15
    with pytest.raises(ValueError):
        ES.apply_strategy(img, "no_such_strategy", {})


def test_ss_strategies_on_colour_corrected_float_image(cuda_device):
    """six_stadigy.py:409-427 called the reference's way: strategyN(color_correction(img)) — the corrected image is
    NOT byte-valued (a channel x 0.85), so this is the primitive route.  Pixels may differ by more than 1/255 only where
    a float difference of ~1e-7 flips a byte in front of CLAHE; those must stay below 1e-3 of the image."""
    frame = synth_image(0, 120, 160)                                     # bluish
    img = frame.astype(np.float32) / 255.0
    kind = compat.detect_image_type(img)
    assert kind == o.ss_image_type(img) == "bluish"
    assert compat.detect_image_type(frame) == kind                       # uint8 frames: histogram route
    cur = compat.color_correction(img, kind)
    assert np.array_equal(cur, o.ss_correct_cast(img, kind))
    assert compat.as_frame(cur) is None
    SS = compat.SSApi
    fns = (SS.strategy1_strong_dehazing, SS.strategy2_medium_dehazing, SS.strategy3_light_dehazing,
           SS.strategy4_clahe_enhancement, SS.strategy5_white_balance, SS.strategy6_histogram_eq)
    for i, fn in enumerate(fns):
        got = fn(cur)
        want = o.ss_strategy(cur, i)
        assert got.dtype == np.float32 and got.shape == cur.shape
        d = np.abs(got.astype(np.float64) - want.astype(np.float64))
        assert (d > 1.0 / 255.0 + 1e-6).mean() < 1e-3, f"strategy {i + 1}: {(d > 1 / 255).sum()} pixels beyond 1/255, max {d.max():.3e}"


# ------------------------------------------------------------------ f-1: file loop with pinned double buffering
def test_process_all_images_all_strategies_streams_and_logs(cuda_device, tmp_path):
    import cv2
    import pandas as pd
    src, dst = tmp_path / "in", tmp_path / "out"
    src.mkdir()
    frames = {}
    for i, (h, w) in enumerate(((96, 128), (96, 128), (96, 128), (80, 100), (80, 100))):
        frames[f"img{i}.png"] = synth_image(30 + i, h, w)
        cv2.imwrite(str(src / f"img{i}.png"), cv2.cvtColor(frames[f"img{i}.png"], cv2.COLOR_RGB2BGR))
    (src / "broken.png").write_bytes(b"not a png")
    log = compat.process_all_images_all_strategies(src, dst, batch=2, workers=2)
    assert len(log) == 5 * 6 and all(r["status"] == "success" for r in log)
    df = pd.read_csv(dst / "processing_log.csv", encoding="utf-8-sig")
    assert list(df.columns) == ["filename", "image_type", "strategy", "strategy_desc", "status", "output_path", "processing_time"]
    assert len(df) == 30
    for name, frame in frames.items():
        ref = o.p6_frame(frame, with_scores=False)
        for s, sname in enumerate(o.SS_NAMES):
            got = cv2.cvtColor(cv2.imread(str(dst / f"{name[:-4]}_{sname}.png")), cv2.COLOR_BGR2RGB)
            d = np.abs(got.astype(np.int16) - ref["out"][s].astype(np.int16))
            assert d.max() <= 1 and (d != 0).mean() < 1e-3, (name, sname)
        assert set(df[df.filename == name].image_type) == {o.IMAGE_TYPES[ref["type"]]}
    compat.close_engines()


# ------------------------------------------------------------------ f-2: build_dataset fed from the device
def test_selfsupervised_build_dataset_and_predict(cuda_device, tmp_path):
    import cv2
    import pandas as pd
    from sklearn.dummy import DummyClassifier
    from sklearn.preprocessing import StandardScaler
    from underwater_image_enhancement_b200.selfsup import SelfSupervisedSystem
    src = tmp_path / "images"
    src.mkdir()
    seeds = list(range(40, 46))
    for s in seeds:
        cv2.imwrite(str(src / f"f{s}.png"), cv2.cvtColor(synth_image(s, 96, 128), cv2.COLOR_RGB2BGR))
    out = tmp_path / "out"

    class Cfg:
        IMAGE_FOLDER, OUTPUT_FOLDER = str(src), str(out)
        STRATEGY_FOLDER, MODEL_FOLDER, REPORT_FOLDER = str(out / "enhanced"), str(out / "models"), str(out / "reports")
        SUPPORTED_FORMATS = [".png"]
        STRATEGIES, QUALITY_WEIGHTS = o.ES_DEFAULT_STRATEGIES, o.ES_QUALITY_WEIGHTS

        @classmethod
        def create_folders(cls):
            for f in (cls.OUTPUT_FOLDER, cls.STRATEGY_FOLDER, cls.MODEL_FOLDER, cls.REPORT_FOLDER):
                os.makedirs(f, exist_ok=True)

    system = SelfSupervisedSystem(Cfg, batch=4, workers=2)
    system.build_dataset()
    assert len(system.dataset) == len(seeds)
    df = pd.read_csv(out / "reports" / "dataset_building.csv")
    names = [p["name"] for p in o.ES_DEFAULT_STRATEGIES.values()]
    assert list(df.columns) == ["filename", "best_strategy", "best_score"] + names        # main.py:167-173
    by_name = {d["filename"]: d for d in system.dataset}
    for s in seeds[:2]:
        ref = o.p5_frame(synth_image(s, 96, 128))
        item = by_name[f"f{s}.png"]
        assert item["best_strategy"] == names[ref["best"]]
        assert abs(item["best_score"] - ref["scores"][ref["best"], 0]) <= 1e-4 * abs(ref["scores"][ref["best"], 0])
        assert item["features"].shape == (79,)
        best_png = cv2.cvtColor(cv2.imread(str(out / "enhanced" / f"f{s}_{item['best_strategy']}.png")), cv2.COLOR_BGR2RGB)
        d = np.abs(best_png.astype(np.int16) - ref["out"][ref["best"]].astype(np.int16))
        assert d.max() <= 1
    X = np.stack([d["features"] for d in system.dataset])
    system.scaler = StandardScaler().fit(X)
    system.classifier = DummyClassifier(strategy="prior").fit(system.scaler.transform(X), [d["best_strategy"] for d in system.dataset])
    pred, proba = system.predict(src / f"f{seeds[0]}.png")
    assert pred in names and abs(sum(proba.values()) - 1.0) < 1e-9
    compat.close_engines()


# ------------------------------------------------------------------ f-3 / f-4: the VGG-guided path
def _reference_vgg():
    sys.path.append(REF)
    import vgg_16_UIE                                     # the reference's own file (network + enhancer)
    return vgg_16_UIE


@needs_ref
def test_enhancement_predictor_batched_vs_reference_predictor(cuda_device, tmp_path):
    """use_trained_model.py's own EnhancementPredictor (baseline/_ref, unmodified, on the same GPU) on a random-init
    checkpoint (seed 0) vs the batched face: parameters and enhanced images."""
    ref_vgg = _reference_vgg()
    from underwater_image_enhancement_b200.predictor import PARAM_ORDER, EnhancementPredictor
    torch.manual_seed(0)
    net = ref_vgg.ImprovedVGGParameterNet(pretrained=False, hidden_dim=256, use_features=True)
    ckpt = tmp_path / "model.pth"
    torch.save({"model_state_dict": net.state_dict()}, ckpt)
    import use_trained_model as ref_utm                   # resolves to baseline/_ref (appended above)
    assert os.path.dirname(os.path.abspath(ref_utm.__file__)) == os.path.abspath(REF)
    ref_pred = ref_utm.EnhancementPredictor(str(ckpt), device="cuda")
    mine = EnhancementPredictor(str(ckpt), device="cuda")
    frames = np.stack([synth_image(70 + i, 270, 480) for i in range(4)])
    enhanced, used = mine.enhance_batch(torch.from_numpy(frames).cuda())
    enhanced, used = enhanced.cpu().numpy(), used.cpu().numpy()
    for i in range(4):
        img = frames[i].astype(np.float32) / 255.0
        p_ref = ref_pred.predict_parameters(img)
        for r, k in enumerate(PARAM_ORDER):
            assert abs(used[r, i] - p_ref[k]) <= 1e-4 * max(1.0, abs(p_ref[k])), (k, used[r, i], p_ref[k])
        e_ref = ref_pred.enhance_image(img, p_ref)
        assert np.abs(enhanced[i].transpose(1, 2, 0) - e_ref).max() < 2e-3
        p_mine = mine.predict_parameters(img)
        assert set(p_mine) == set(p_ref)
        e_mine = mine.enhance_image(img, p_ref)            # same parameters -> the enhancer alone
        assert np.abs(e_mine - e_ref).max() < 1e-5


@needs_ref
def test_autograd_enhancer_matches_torch_autograd(cuda_device):
    """vgg_16_UIE.py:529 / deep_learning_parameters.py:281: gradients of a loss through the enhancer with respect to
    omega and gamma, against torch autograd on the reference's own DifferentiableEnhancement."""
    ref_vgg = _reference_vgg()
    from underwater_image_enhancement_b200.vgg import DifferentiableEnhancement
    torch.manual_seed(1)
    B, H, W = 3, 90, 120
    frames = np.stack([synth_image(80 + i, H, W) for i in range(B)])
    img = (torch.from_numpy(frames).cuda().float() / 255.0).permute(0, 3, 1, 2).contiguous()
    weight = torch.rand_like(img)
    base = {"L_low": torch.tensor([[5.0], [10.0], [2.0]], device="cuda"), "L_high": torch.tensor([[95.0], [90.0], [98.0]], device="cuda")}

    def run(module, om, ga):
        p = dict(base, omega=om, gamma=ga)
        out = module(img, p)
        (out * weight).sum().backward()
        return out.detach()

    om_r = torch.tensor([[0.5], [0.7], [0.35]], device="cuda", requires_grad=True)
    ga_r = torch.tensor([[1.2], [0.9], [1.4]], device="cuda", requires_grad=True)
    out_ref = run(ref_vgg.DifferentiableEnhancement().cuda(), om_r, ga_r)
    om_m = om_r.detach().clone().requires_grad_(True)
    ga_m = ga_r.detach().clone().requires_grad_(True)
    out_mine = run(DifferentiableEnhancement(), om_m, ga_m)
    assert (out_mine - out_ref).abs().max() < 1e-5
    assert om_m.grad is not None and ga_m.grad is not None
    assert torch.allclose(om_m.grad, om_r.grad, rtol=2e-3, atol=1e-2), (om_m.grad, om_r.grad)
    assert torch.allclose(ga_m.grad, ga_r.grad, rtol=2e-3, atol=1e-2), (ga_m.grad, ga_r.grad)
    with torch.no_grad():                                   # inference path: no graph, same numbers
        out_inf = DifferentiableEnhancement()(img, dict(base, omega=om_m, gamma=ga_m))
    assert not out_inf.requires_grad and (out_inf - out_ref).abs().max() < 1e-5
