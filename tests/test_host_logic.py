"""Host-side logic that needs no GPU: byte-valued image classification, the pinned double-buffered frame reader / writer
(plain host memory without a device), the bench's workload naming and NUMA helpers, the shim modules' surface."""
from __future__ import annotations

import ast
import os
import sys

import cv2
import numpy as np

from underwater_image_enhancement_b200 import compat, frame_io
from underwater_image_enhancement_b200.synth import synth_image

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_as_frame_tells_bytes_from_floats():
    f = synth_image(1, 40, 50)
    assert np.array_equal(compat.as_frame(f), f)
    assert np.array_equal(compat.as_frame(f.astype(np.float32) / 255.0), f)          # main.py:108
    assert np.array_equal(compat.as_frame(f.astype(np.float64) / 255.0), f)
    rng = np.random.default_rng(0)
    assert compat.as_frame(rng.random((40, 50, 3))) is None                          # example_usage.py:27
    assert compat.as_frame((f.astype(np.float32) / 255.0) * np.float32(0.85)) is None  # six_stadigy.py:309
    assert compat.as_frame(np.ones((8, 8, 3)) * 0.5) is None
    assert compat.as_frame(np.ones((8, 8, 3)) * 1.5) is None
    assert np.array_equal(compat.as_frame(np.random.default_rng(1).choice([0.0, 1.0], size=(8, 8, 3))) % 255, np.zeros((8, 8, 3)))


def test_frame_reader_batches_by_size_and_reports_failures(tmp_path):
    sizes = [(32, 48)] * 3 + [(40, 40)] * 2 + [(32, 48)]
    frames = []
    for i, (h, w) in enumerate(sizes):
        fr = synth_image(i, h, w)
        frames.append(fr)
        cv2.imwrite(str(tmp_path / f"a{i:02d}.png"), cv2.cvtColor(fr, cv2.COLOR_RGB2BGR))
    (tmp_path / "a99.png").write_bytes(b"junk")
    files = frame_io.list_images(tmp_path)
    assert [p.name for p in files] == [f"a{i:02d}.png" for i in range(6)] + ["a99.png"]
    reader = frame_io.FrameReader(files, batch=2, workers=2)
    got = []
    for paths, batch in reader:
        assert batch.dtype == np.uint8 and batch.shape[0] == len(paths) <= 2
        assert len({b.shape for b in batch}) == 1
        got += [(p.name, b.copy()) for p, b in zip(paths, batch)]
    reader.close()
    assert [n for n, _ in got] == [f"a{i:02d}.png" for i in range(6)]                # order kept, sizes never mixed
    for (_, b), fr in zip(got, frames):
        assert np.array_equal(b, fr)
    assert [p.name for p in reader.failed] == ["a99.png"]


def test_frame_writer_slots_and_files(tmp_path):
    w = frame_io.FrameWriter(workers=2)
    a = w.slot((2, 16, 16, 3))
    a[:] = 7
    f1 = w.write(tmp_path / "x.png", a[0])
    b = w.slot((2, 16, 16, 3))
    assert b.ctypes.data != a.ctypes.data                                            # double buffering
    c = w.slot((2, 16, 16, 3))                                                       # back to the first slot: its write is done
    assert c.ctypes.data == a.ctypes.data and f1.done() and f1.result() is True
    w.close()
    assert np.all(cv2.imread(str(tmp_path / "x.png")) == 7)


def test_bench_arms_share_one_workload_name_and_parse_cpulists():
    sys.path.insert(0, ROOT)
    import bench
    assert bench.workload_name("six", 256) == bench.workload_name("six", 256)
    src = open(os.path.join(ROOT, "bench.py")).read()
    assert src.count('"workload": workload_name(args.library') == 2                 # native and reference arm
    assert bench._parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    info = bench.bind_to_gpu_numa.__doc__
    assert "numa_node" in info and "topo" in info


def test_shim_modules_export_the_reference_surface():
    """Names the reference's callers import (main.py:22-25, example_usage.py:15-18, use_trained_model.py:9)."""
    want = {"enhancement_strategies.py": {"EnhancementStrategies", "test_strategies"},
            "six_stadigy.py": {"EnhancementStrategies", "detect_image_type", "color_correction", "process_all_images_all_strategies"},
            "quality_assessment.py": {"QualityAssessment", "test_quality_assessment"},
            "feature_extraction.py": {"FeatureExtractor", "test_feature_extraction"},
            "use_trained_model.py": {"EnhancementPredictor"}}
    for name, names in want.items():
        tree = ast.parse(open(os.path.join(ROOT, "shims", name)).read())
        have = set()
        for node in ast.walk(tree):
            if isinstance(node, ast.FunctionDef):
                have.add(node.name)
            elif isinstance(node, ast.ImportFrom):
                have |= {a.asname or a.name for a in node.names}
        assert names <= have, (name, names - have)
    for cls, methods in ((compat.ESApi, ("guided_filter", "estimate_atmospheric_light", "compute_Q", "get_brightest_pixel",
                                         "estimate_transmission", "recover_image", "color_enhancement", "gamma_correction",
                                         "clahe_enhancement", "histogram_equalization", "apply_strong_dehazing",
                                         "apply_medium_dehazing", "apply_clahe_enhancement", "apply_light_enhancement",
                                         "apply_histogram_equalization", "apply_strategy")),
                         (compat.SSApi, ("guided_filter", "estimate_atmospheric_light", "compute_Q", "get_brightest_pixel",
                                         "estimate_transmission", "restore_image", "enhance_contrast", "apply_clahe",
                                         "white_balance", "gamma_correction", "strategy1_strong_dehazing",
                                         "strategy2_medium_dehazing", "strategy3_light_dehazing", "strategy4_clahe_enhancement",
                                         "strategy5_white_balance", "strategy6_histogram_eq"))):
        for m in methods:
            assert callable(getattr(cls, m)), (cls.__name__, m)


def test_unknown_strategy_raises_without_a_device():
    import pytest
    with pytest.raises(ValueError):
        compat.ESApi.apply_strategy(np.zeros((4, 4, 3)), "nope", {})
