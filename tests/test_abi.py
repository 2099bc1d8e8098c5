"""The C-ABI library loads without a GPU and exports every symbol include/uwie.h declares; the
argument checks and presets work without a device; compute entry points fail LOUDLY (no CPU
fallback) when there is no CUDA device."""
from __future__ import annotations

import ctypes as C
import os
import re

import pytest
import torch

from underwater_image_enhancement_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "uwie.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(uwie_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree():
    assert _declared_symbols() == sorted(_lib.EXPORTS)


def test_library_exports_every_declared_symbol(native_lib):
    for name in _declared_symbols():
        assert hasattr(native_lib, name), name
    assert native_lib.uwie_version() == 100


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "underwater_image_enhancement_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dirpath, f), errors="replace").read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f
                assert "restatement" not in src, f


def test_presets(native_lib):
    lib, st = _lib.preset(_lib.LIB_SIX)
    assert len(st) == 6 and lib.f64 == 0 and lib.colour_cast == 1
    assert [s.dehaze for s in st] == [1, 1, 1, 0, 0, 0]
    assert (st[0].omega, st[0].radius, st[0].eps) == (0.3, 20, 0.5)           # SS:233-236
    assert [st[3].post_lo[1], st[3].post_hi[1]] == [3.0, 97.0]                 # white_balance(3), SS:266
    assert abs(sum(lib.weights) - 1.0) < 1e-12                                 # QA:229-238
    lib, st = _lib.preset(_lib.LIB_ES)
    assert len(st) == 5 and lib.f64 == 1 and lib.gamma_inverse == 1
    assert st[4].equalize == 1 and st[3].clahe == 1
    assert list(lib.weights)[6:] == [0.0, 0.0]                                 # config.py:78-85 has 6 keys
    with pytest.raises(_lib.UwieError):
        _lib.preset(7)


def test_workspace_and_argument_checks(native_lib):
    lib, st = _lib.preset(_lib.LIB_SIX)
    arr = (_lib.Strategy * len(st))(*st)
    need = native_lib.uwie_workspace_bytes(C.byref(lib), arr, len(st), 4, 1080, 1920)
    assert need > 4 * 6 * 1080 * 1920 * 3
    assert native_lib.uwie_workspace_bytes(C.byref(lib), arr, len(st), 0, 1080, 1920) == 0
    assert native_lib.uwie_workspace_bytes(C.byref(lib), arr, 7, 1, 64, 64) == 0
    bad = (_lib.Strategy * 1)(st[0])
    bad[0].radius = 0
    assert native_lib.uwie_workspace_bytes(C.byref(lib), bad, 1, 1, 64, 64) == 0
    assert b"radius" in native_lib.uwie_last_error()


@pytest.mark.skipif(torch.cuda.is_available(), reason="only meaningful without a CUDA device")
def test_no_cpu_fallback(native_lib):
    lib, st = _lib.preset(_lib.LIB_SIX)
    arr = (_lib.Strategy * len(st))(*st)
    h = C.c_void_p()
    rc = native_lib.uwie_engine_create(C.byref(h), C.byref(lib), arr, len(st), 1, 64, 64, None, 0)
    assert rc == _lib.ERR_CUDA and not h.value
    assert b"no CPU path" in native_lib.uwie_last_error()
    from underwater_image_enhancement_b200.engine import Engine
    with pytest.raises(RuntimeError):
        Engine("six", 64, 64)


def test_every_entry_point_cites_the_reference_or_says_it_has_no_counterpart():
    """include/uwie.h: the comment block above each group of declarations names the reference code the entry points replace
    (file:line, or the header's own abbreviations ES: / SS: / QA: / FE: / VGG: / UTM:), or says that there is none."""
    src = open(os.path.join(ROOT, "include", "uwie.h")).read()
    cite = re.compile(r"(\w+\.py:\d+|\b(ES|SS|QA|FE|VGG|UTM|DLP|MAIN):\d+|NO counterpart|no counterpart)")
    housekeeping = {"uwie_version", "uwie_last_error", "uwie_workspace_bytes", "uwie_engine_create", "uwie_engine_destroy",
                    "uwie_engine_check", "uwie_engine_last_launches", "uwie_engine_graph_replays", "uwie_engine_debug",
                    "uwie_engine_profile", "uwie_engine_profile_read", "uwie_selftest_division", "uwie_engine_host_prepare"}
    missing = []
    for m in re.finditer(r"^[a-z_0-9 \*]+?\b(uwie_[a-z0-9_]+)\s*\(", src, flags=re.M):
        name = m.group(1)
        if name in housekeeping or name.endswith("_workspace_bytes"):
            continue
        # the nearest comment block that ends before this declaration
        end = src.rfind("*/", 0, m.start())
        start = src.rfind("/*", 0, end)
        assert start >= 0 and end > start, name
        if not cite.search(src[start:end]):
            missing.append(name)
    assert not missing, missing
