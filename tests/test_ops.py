"""torch.ops.uwie.*: registration (CPU box) and parity through the operators (GPU box)."""
from __future__ import annotations

import numpy as np
import pytest
import torch

from conftest import golden_paths, load_golden
from oracle import restatement as o
from underwater_image_enhancement_b200 import ops  # noqa: F401  (registers the operators)
from underwater_image_enhancement_b200.synth import synth_image

OPS = ("enhance_score_p6", "enhance_score_p5", "quality_scores", "features79", "vgg_enhance", "vgg_features",
       "vgg_preprocess", "clahe_lab", "equalize_hist", "transmission_head", "atmospheric_light", "transmission", "recover",
       "percentile_stretch", "gamma")


def test_operators_are_registered_for_cuda_only():
    for name in OPS:
        assert hasattr(torch.ops.uwie, name), name
    frames = torch.zeros((1, 8, 8, 3), dtype=torch.uint8)
    with pytest.raises((NotImplementedError, RuntimeError)):        # no CPU kernel: the dispatcher refuses
        torch.ops.uwie.enhance_score_p6(frames)
    with pytest.raises((NotImplementedError, RuntimeError)):
        torch.ops.uwie.features79(frames)


@pytest.mark.gpu
def test_fused_operators_match_reference_fixture():
    g = load_golden(golden_paths()[0])
    frames = torch.from_numpy(np.stack([g["frame"]] * 3)).cuda()     # 3 frames: more than one would fit a 1-frame engine
    out, scores, best, kind = torch.ops.uwie.enhance_score_p6(frames)
    for i in range(3):
        assert np.array_equal(out[i].cpu().numpy(), g["p6_out"])
        assert np.allclose(scores[i].cpu().numpy(), g["p6_scores"], rtol=1e-4)
    assert kind.cpu().tolist() == [int(g["p6_type"])] * 3
    assert best.cpu().tolist() == [int(np.argmax(g["p6_scores"][:, 0]))] * 3
    best_out, scores5, best5 = torch.ops.uwie.enhance_score_p5(frames[:1])
    want = int(np.argmax(g["p5_scores"][:, 0]))
    assert int(best5[0]) == want
    assert np.array_equal(best_out[0].cpu().numpy(), g["p5_out"][want])
    assert np.allclose(scores5[0].cpu().numpy(), g["p5_scores"], rtol=1e-4)
    ops.close_engines()


@pytest.mark.gpu
def test_other_operators_match_oracle():
    frame = synth_image(41, 90, 122)
    frames = torch.from_numpy(frame[None]).cuda()
    feats = torch.ops.uwie.features79(frames)[0].cpu().numpy()
    want = o.fe_all(frame)
    assert np.allclose(feats, want, rtol=1e-4, atol=1e-6)
    img = frame.astype(np.float32) / 255.0
    w = [0.20, 0.20, 0.15, 0.15, 0.10, 0.10, 0.05, 0.05]
    row = torch.ops.uwie.quality_scores(torch.from_numpy(img[None]).cuda(), w)[0].cpu().numpy()
    total, sub = o.qa_total(img)
    assert np.allclose(row, o.scores_row(total, sub), rtol=1e-4)
    nchw = torch.from_numpy(img.transpose(2, 0, 1)[None].copy()).cuda()
    lo, hi, om = torch.tensor([[5.0]]), torch.tensor([[95.0]]), torch.tensor([[0.6]])
    enh = torch.ops.uwie.vgg_enhance(nchw, lo, hi, om, None).cpu().numpy()
    assert np.array_equal(enh, o.vgg_enhance(nchw.cpu().numpy(), [5.0], [95.0], [0.6], None))
    assert np.allclose(torch.ops.uwie.vgg_features(frames)[0].cpu().numpy(), o.vgg_features(frame), rtol=1e-5, atol=1e-7)
    assert np.array_equal(torch.ops.uwie.vgg_preprocess(frames, 224)[0].cpu().numpy(), o.vgg_preprocess(frame))


@pytest.mark.gpu
def test_primitive_operators_match_cv2_and_oracle():
    import cv2
    frame = synth_image(42, 96, 128)
    L = torch.from_numpy(np.ascontiguousarray(cv2.cvtColor(frame, cv2.COLOR_RGB2LAB)[:, :, 0])[None]).cuda()
    got = torch.ops.uwie.clahe_lab(L, 2.0, 8, 8)[0].cpu().numpy()
    assert np.array_equal(got, cv2.createCLAHE(clipLimit=2.0, tileGridSize=(8, 8)).apply(L[0].cpu().numpy()))
    eq = torch.ops.uwie.equalize_hist(torch.from_numpy(frame[None]).cuda())[0].cpu().numpy()
    assert np.array_equal(eq[:, :, 1], cv2.equalizeHist(np.ascontiguousarray(frame[:, :, 1])))
    img = np.random.default_rng(3).random((96, 128, 3))
    d = torch.from_numpy(img).cuda()
    A = torch.ops.uwie.atmospheric_light(d)
    assert np.array_equal(A.cpu().numpy(), o.quadtree_airlight(img, 1))
    t = torch.ops.uwie.transmission(d, A, 0.5, 15, 1e-3, 1e-10, False)
    plane = np.tile(A.cpu().numpy().reshape(1, 1, 3), (96, 128, 1))
    assert np.abs(t.cpu().numpy() - o.es_transmission(img, plane, 0.5, 15)).max() < 1e-9
    rec = torch.ops.uwie.recover(d, t, A, True)
    st = torch.ops.uwie.percentile_stretch(rec, 10.0, 95.0, 1e-10)
    assert np.array_equal(st.cpu().numpy(), o.percentile_stretch(rec.cpu().numpy(), 10, 95, 1e-10))
    g = torch.ops.uwie.gamma(st, 1 / 1.2, True)
    assert np.abs(g.cpu().numpy() - o.es_gamma(st.cpu().numpy(), 1.2)).max() < 1e-15
    th = torch.ops.uwie.transmission_head(torch.from_numpy(frame[None]).cuda(), torch.tensor([[0.3, 0.6, 0.7]], device="cuda"), 0.5, 1e-6, True)
    assert th.shape == (1, 96, 128) and th.dtype == torch.float32
