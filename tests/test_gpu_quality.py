"""Stand-alone scoring (uwie_quality_scores) against the oracle and the reference fixtures: the enhanced
float images of the fixtures are scored again and must reproduce the reference's scores."""
from __future__ import annotations

import numpy as np
import pytest

from conftest import golden_paths, load_golden
from oracle import restatement as o
from underwater_image_enhancement_b200.engine import QA_KEYS
from underwater_image_enhancement_b200.quality import QualityAssessment, quality_scores_batch
from underwater_image_enhancement_b200.synth import synth_image

pytestmark = pytest.mark.gpu
RTOL = 1e-4      # north_star: 1e-4 relative on quality scores


def _rel(a, b):
    return np.abs(a - b) / np.maximum(np.abs(b), 1e-9)


def test_scores_of_reference_outputs():
    for path in golden_paths():
        g = load_golden(path)
        # the stored uint8 outputs, scored as the float32 image bytes / 255 (the reference's own totals
        # were taken on the float outputs, which differ from bytes / 255 wherever a gamma came last)
        sel = [0, 1, 3, 4, 5]
        got = quality_scores_batch(g["p6_out"][sel]).cpu().numpy()
        for i, s in enumerate(sel):
            total, sub = o.qa_total(g["p6_out"][s].astype(np.float32) / 255.0)
            assert _rel(got[i], o.scores_row(total, sub)).max() <= RTOL, (path, s)
            cols = [1, 2, 3, 4, 5, 6, 8]                 # everything but colourfulness is a function of the bytes
            assert _rel(got[i][cols], g["p6_scores"][s][cols]).max() <= RTOL, (path, s)
        if "p6_s3_float" in g:                           # genuinely float32 image
            got = quality_scores_batch(g["p6_s3_float"]).cpu().numpy()[0]
            assert _rel(got, g["p6_scores"][2]).max() <= RTOL
        if "p5_strong_float" in g:                       # float64 image, config.py weights
            w = {"contrast": 0.25, "sharpness": 0.20, "entropy": 0.15, "saturation": 0.15, "brightness": 0.15,
                 "edge_density": 0.10}
            got = quality_scores_batch(g["p5_strong_float"], weights=w).cpu().numpy()[0]
            assert _rel(got, g["p5_scores"][0]).max() <= RTOL


@pytest.mark.parametrize("shape", [(33, 250), (720, 1280)])
def test_scores_match_oracle(shape):
    h, w = shape
    frame = synth_image(31, h, w)
    img = frame.astype(np.float32) / 255.0
    total, sub = QualityAssessment.comprehensive_assessment(img)
    ref_total, ref_sub = o.qa_total(img)
    assert abs(total - ref_total) <= RTOL * abs(ref_total)
    for k in QA_KEYS:
        assert abs(sub[k] - ref_sub[k]) <= RTOL * max(abs(ref_sub[k]), 1e-9), k
    assert sub["sharpness"] == 50.0                       # QA:49, 250-251
    assert QualityAssessment.assess_contrast(img) == sub["contrast"]


def test_constant_images():
    """quality_assessment.py:298 style inputs: constant images."""
    for v in (0.0, 0.5, 1.0):
        img = np.full((40, 56, 3), v, np.float32)
        total, sub = QualityAssessment.comprehensive_assessment(img)
        ref_total, ref_sub = o.qa_total(img)
        assert abs(total - ref_total) <= 1e-4 * max(abs(ref_total), 1.0)
        for k in QA_KEYS:
            assert abs(sub[k] - ref_sub[k]) <= 1e-4 * max(abs(ref_sub[k]), 1.0), (v, k)
