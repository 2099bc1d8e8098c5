"""Drop-in for the reference's feature_extraction.py (feature_extraction.py:13-324) — on the B200."""
import numpy as np

import _bootstrap  # noqa: F401
from underwater_image_enhancement_b200.features import FeatureExtractor  # noqa: F401


def test_feature_extraction():
    """feature_extraction.py:300-320."""
    print("測試特徵提取...")
    test_img = np.random.rand(256, 256, 3)
    features = FeatureExtractor.extract_all_features(test_img)
    print(f"特徵維度: {len(features)}")
    print(f"特徵範圍: [{np.min(features):.4f}, {np.max(features):.4f}]")
    print(f"特徵均值: {np.mean(features):.4f}")
    print(f"特徵標準差: {np.std(features):.4f}")
    if np.any(np.isnan(features)):
        print("警告: 特徵中包含 NaN")
    if np.any(np.isinf(features)):
        print("警告: 特徵中包含 Inf")
    print("✓ 特徵提取測試完成")
    return features


if __name__ == '__main__':
    test_feature_extraction()
