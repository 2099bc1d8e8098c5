"""Drop-in for the reference's six_stadigy.py (six_stadigy.py:22-470): EnhancementStrategies.strategy1..6 and their
primitives, detect_image_type, color_correction, process_all_images_all_strategies — on the B200."""
import _bootstrap  # noqa: F401
from underwater_image_enhancement_b200.compat import SSApi as EnhancementStrategies  # noqa: F401
from underwater_image_enhancement_b200.compat import (color_correction, detect_image_type,  # noqa: F401
                                                      process_all_images_all_strategies)

if __name__ == "__main__":
    import os
    import sys
    if len(sys.argv) != 3 or not os.path.exists(sys.argv[1]):
        print("usage: python six_stadigy.py INPUT_FOLDER OUTPUT_FOLDER")
    else:
        process_all_images_all_strategies(sys.argv[1], sys.argv[2])
