"""Drop-in for the reference's quality_assessment.py (quality_assessment.py:11-322) — on the B200."""
import numpy as np

import _bootstrap  # noqa: F401
from underwater_image_enhancement_b200.quality import QualityAssessment  # noqa: F401


def test_quality_assessment():
    """quality_assessment.py:289-318."""
    print("測試品質評估...")
    test_images = {
        'Random': np.random.rand(256, 256, 3),
        'Low Contrast': np.ones((256, 256, 3)) * 0.5,
        'High Contrast': np.random.choice([0.0, 1.0], size=(256, 256, 3)),
        'Dark': np.random.rand(256, 256, 3) * 0.3,
        'Bright': 0.7 + np.random.rand(256, 256, 3) * 0.3
    }
    print("\n評估結果:")
    print("-" * 80)
    print(f"{'影像':<15} {'總分':<10} {'對比度':<10} {'清晰度':<10} {'熵':<10} {'飽和度':<10}")
    print("-" * 80)
    out = {}
    for name, img in test_images.items():
        total_score, scores = QualityAssessment.comprehensive_assessment(img)
        out[name] = (total_score, scores)
        print(f"{name:<15} {total_score:>6.2f}    {scores['contrast']:>6.2f}    {scores['sharpness']:>6.2f}    "
              f"{scores['entropy']:>6.2f}    {scores['saturation']:>6.2f}")
    print("-" * 80)
    print("✓ 品質評估測試完成")
    return out


if __name__ == '__main__':
    test_quality_assessment()
