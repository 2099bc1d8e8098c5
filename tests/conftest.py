"""Shared fixtures.  `-m "not gpu"` runs on the build container (no CUDA device); `-m gpu` runs on
the B200 box and goes through the C ABI of libuwie_b200.so."""
from __future__ import annotations

import glob
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def golden_paths():
    return sorted(glob.glob(os.path.join(GOLDEN_DIR, "ref_seed*.npz")))


def golden_ids():
    return [os.path.basename(p)[4:-4] for p in golden_paths()]


@pytest.fixture(scope="session")
def native_lib():
    """Builds (if needed) and loads the product library; never falls back to anything else."""
    from underwater_image_enhancement_b200 import build as _build
    _build.build()
    from underwater_image_enhancement_b200 import _lib
    return _lib.load()


@pytest.fixture(scope="session")
def cuda_device():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("gpu-marked test started without a CUDA device")
    return torch.device("cuda:0")


def load_golden(path):
    with np.load(path) as z:
        return {k: z[k] for k in z.files}
