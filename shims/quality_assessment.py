"""Drop-in for the reference's quality_assessment.py (quality_assessment.py:11-322) — on the B200."""
import numpy as np

import _bootstrap  # noqa: F401
from underwater_image_enhancement_b200.quality import QualityAssessment  # noqa: F401

_COLUMNS = (("影像", None), ("總分", None), ("對比度", "contrast"), ("清晰度", "sharpness"), ("熵", "entropy"), ("飽和度", "saturation"))


def test_quality_assessment(size=256, seed=None):
    """Module self-test with the five synthetic images of quality_assessment.py:289-318 (random, flat, binary, dark,
    bright); prints the same table and returns {name: (total, sub-scores)}."""
    rng = np.random.default_rng(seed)
    shape = (size, size, 3)
    cases = {"Random": rng.random(shape), "Low Contrast": np.full(shape, 0.5),
             "High Contrast": rng.integers(0, 2, shape).astype(np.float64),
             "Dark": 0.3 * rng.random(shape), "Bright": 0.7 + 0.3 * rng.random(shape)}
    rule = "-" * 80
    print("測試品質評估...\n\n評估結果:\n" + rule)
    print(f"{_COLUMNS[0][0]:<15} " + " ".join(f"{title:<10}" for title, _ in _COLUMNS[1:]) + "\n" + rule)
    results = {}
    for label, image in cases.items():
        results[label] = QualityAssessment.comprehensive_assessment(image)
        total, sub = results[label]
        cells = [total] + [sub[key] for _, key in _COLUMNS[2:]]
        print(f"{label:<15} " + "    ".join(f"{v:>6.2f}" for v in cells))
    print(rule + "\n✓ 品質評估測試完成")
    return results


if __name__ == '__main__':
    test_quality_assessment()
