"""Drop-in for the reference's feature_extraction.py (feature_extraction.py:13-324) — on the B200."""
import numpy as np

import _bootstrap  # noqa: F401
from underwater_image_enhancement_b200.features import FeatureExtractor  # noqa: F401


def test_feature_extraction(size=256, seed=None):
    """Module self-test of feature_extraction.py:300-320: the 79-vector of one random image, with the same report."""
    vec = FeatureExtractor.extract_all_features(np.random.default_rng(seed).random((size, size, 3)))
    report = (("特徵維度", f"{len(vec)}"), ("特徵範圍", f"[{vec.min():.4f}, {vec.max():.4f}]"),
              ("特徵均值", f"{vec.mean():.4f}"), ("特徵標準差", f"{vec.std():.4f}"))
    print("測試特徵提取...")
    for label, text in report:
        print(f"{label}: {text}")
    for bad, what in ((np.isnan(vec), "NaN"), (np.isinf(vec), "Inf")):
        if bad.any():
            print(f"警告: 特徵中包含 {what}")
    print("✓ 特徵提取測試完成")
    return vec


if __name__ == '__main__':
    test_feature_extraction()
