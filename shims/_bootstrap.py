"""Makes the B200-native package importable from the shim modules: either it is already importable (installed, or its
repository root is on PYTHONPATH / UWIE_B200_ROOT), or the shims still live in <repo>/shims/ and the package is one
level up."""
import importlib.util
import os
import sys

if importlib.util.find_spec("underwater_image_enhancement_b200") is None:
    for root in (os.environ.get("UWIE_B200_ROOT", ""), os.path.dirname(os.path.dirname(os.path.abspath(__file__)))):
        if root and os.path.isdir(os.path.join(root, "underwater_image_enhancement_b200")):
            sys.path.append(root)
            break
