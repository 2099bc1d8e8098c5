"""The oracle against the UNMODIFIED reference at the bench size (1920x1080), on the CPU: the frames of
tests/test_gpu_parity_1080p.py are the ones the CUDA path is compared with the oracle on, so oracle == reference here
closes the chain CUDA == oracle == reference at full size.  Needs the reference (baseline/_ref staged by
__graft_entry__.build(), or /root/reference); skipped where neither exists.  tools/oracle_vs_reference_1080p.py runs all
five frames and writes profiles/r02_oracle_vs_reference_1080p.json."""
from __future__ import annotations

import numpy as np
import pytest

from oracle import refload, restatement as o
from underwater_image_enhancement_b200.synth import synth_image

pytestmark = pytest.mark.skipif(refload.reference_dir() is None, reason="the reference tree is not available here")


@pytest.mark.parametrize("library,seed", [("six", 1), ("es", 0)])
def test_oracle_equals_reference_at_1080p(library, seed):
    mods = refload.load_reference()
    frame = synth_image(seed, 1080, 1920)
    ref = refload.p6_frame(mods, frame) if library == "six" else refload.p5_frame(mods, frame)
    ora = o.p6_frame(frame) if library == "six" else o.p5_frame(frame)
    assert np.array_equal(ref["out"], ora["out"])
    assert np.array_equal(ref["scores"], ora["scores"])
    if library == "es":
        assert ref["best"] == ora["best"]
