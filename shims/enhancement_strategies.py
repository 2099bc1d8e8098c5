"""Drop-in for the reference's enhancement_strategies.py (same module name, class, static methods, defaults, return
types and error behaviour; enhancement_strategies.py:13-543) — every operation runs on the B200 through
libuwie_b200.so.  Put this directory first on sys.path and main.py / example_usage.py run unchanged."""
import numpy as np

import _bootstrap  # noqa: F401
from underwater_image_enhancement_b200.compat import ESApi as EnhancementStrategies  # noqa: F401


def test_strategies():
    """enhancement_strategies.py:511-540 (the figure is written only when matplotlib is installed)."""
    test_img = np.random.rand(256, 256, 3)
    from config import Config
    results = {}
    for strategy_name, params in Config.STRATEGIES.items():
        print(f"測試策略: {params['name']}")
        try:
            results[params['name']] = EnhancementStrategies.apply_strategy(test_img, strategy_name, params)
            print("  ✓ 成功")
        except Exception as e:                                  # noqa: BLE001
            print(f"  ✗ 失敗: {e}")
    try:
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
        fig, axes = plt.subplots(2, 3, figsize=(15, 10))
        for ax, (name, img) in zip(axes.flatten(), results.items()):
            ax.imshow(img); ax.set_title(name); ax.axis('off')
        plt.tight_layout()
        plt.savefig('strategy_test.png', dpi=150)
        print("\n測試完成，結果已儲存為 strategy_test.png")
    except ImportError:
        print("\n測試完成")
    return results


if __name__ == '__main__':
    print("測試增強策略...")
    test_strategies()
