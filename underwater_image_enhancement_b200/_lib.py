"""ctypes binding of include/uwie.h.  Fails loudly when the native library is missing: there is no
Python or CPU fallback for the hot path."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libuwie_b200.so")

MAX_STRATEGIES = 6
SCORE_COLS = 9
LIB_SIX, LIB_ES = 0, 1
OK, ERR_ARG, ERR_CUDA, ERR_NOMEM, ERR_SELECT = 0, -1, -2, -3, -4

# every symbol include/uwie.h declares (tests check that the library exports all of them)
EXPORTS = ("uwie_version", "uwie_last_error", "uwie_preset", "uwie_workspace_bytes", "uwie_engine_create",
           "uwie_engine_destroy", "uwie_enhance_score", "uwie_engine_host_prepare", "uwie_enhance_score_host",
           "uwie_engine_image_types", "uwie_engine_check",
           "uwie_engine_last_launches", "uwie_engine_graph_replays", "uwie_engine_debug", "uwie_engine_profile",
           "uwie_engine_profile_read", "uwie_vgg_workspace_bytes", "uwie_vgg_enhance", "uwie_vgg_enhance_dev", "uwie_vgg_enhance_u8", "uwie_vgg_enhance_bwd", "uwie_vgg_features_u8",
           "uwie_vgg_preprocess_u8", "uwie_vgg_preprocess_f32", "uwie_features79_workspace_bytes",
           "uwie_features79", "uwie_quality_workspace_bytes", "uwie_quality_scores", "uwie_selftest_division",
           "uwie_prim_transmission_head_u8", "uwie_prim_clahe_lab_workspace_bytes", "uwie_prim_clahe_lab",
           "uwie_prim_equalize_hist", "uwie_prim_equalize_hist_static", "uwie_prim_quantise", "uwie_prim_atmospheric_light_workspace_bytes",
           "uwie_prim_atmospheric_light", "uwie_prim_guided_filter_workspace_bytes", "uwie_prim_guided_filter",
           "uwie_prim_transmission_workspace_bytes", "uwie_prim_transmission", "uwie_prim_recover",
           "uwie_prim_percentile_stretch_workspace_bytes", "uwie_prim_percentile_stretch", "uwie_prim_gamma",
           "uwie_prim_scale_clip", "uwie_prim_channel_means", "uwie_prim_region_score_workspace_bytes",
           "uwie_prim_region_score", "uwie_prim_brightest_pixel")
FEATURES = 79
VGG_FEATURES = 79
STAGE_NAME_LEN = 32


class Strategy(C.Structure):
    _fields_ = [("dehaze", C.c_int), ("omega", C.c_double), ("radius", C.c_int), ("eps", C.c_double),
                ("n_pre", C.c_int), ("pre_lo", C.c_double * 2), ("pre_hi", C.c_double * 2),
                ("equalize", C.c_int), ("clahe", C.c_int), ("clip_limit", C.c_double),
                ("tiles_x", C.c_int), ("tiles_y", C.c_int),
                ("n_post", C.c_int), ("post_lo", C.c_double * 2), ("post_hi", C.c_double * 2),
                ("gamma_on", C.c_int), ("gamma", C.c_double)]


class Library(C.Structure):
    _fields_ = [("f64", C.c_int), ("colour_cast", C.c_int), ("dark_guard", C.c_double),
                ("clip_t0", C.c_int), ("stretch_guard", C.c_double), ("gamma_inverse", C.c_int),
                ("weights", C.c_double * 8)]


class UwieError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"uwie error {code}: {msg}")
        self.code = code


_lib = None


def load() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -m underwater_image_enhancement_b200.build` "
            "(nvcc, sm_100a).  The hot path has no CPU or pure-Python fallback.")
    lib = C.CDLL(LIB_PATH)
    lib.uwie_version.restype = C.c_int
    lib.uwie_last_error.restype = C.c_char_p
    lib.uwie_preset.argtypes = [C.c_int, C.POINTER(Library), C.POINTER(Strategy)]
    lib.uwie_preset.restype = C.c_int
    lib.uwie_workspace_bytes.argtypes = [C.POINTER(Library), C.POINTER(Strategy), C.c_int, C.c_int, C.c_int, C.c_int]
    lib.uwie_workspace_bytes.restype = C.c_size_t
    lib.uwie_engine_create.argtypes = [C.POINTER(C.c_void_p), C.POINTER(Library), C.POINTER(Strategy), C.c_int,
                                       C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t]
    lib.uwie_engine_create.restype = C.c_int
    lib.uwie_engine_destroy.argtypes = [C.c_void_p]
    lib.uwie_engine_destroy.restype = None
    lib.uwie_enhance_score.argtypes = [C.c_void_p, C.c_void_p, C.c_int] + [C.c_void_p] * 7
    lib.uwie_enhance_score.restype = C.c_int
    lib.uwie_enhance_score_host.argtypes = [C.c_void_p, C.c_void_p, C.c_int] + [C.c_void_p] * 7
    lib.uwie_enhance_score_host.restype = C.c_int
    lib.uwie_engine_image_types.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
    lib.uwie_engine_image_types.restype = C.c_int
    lib.uwie_engine_host_prepare.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
    lib.uwie_engine_host_prepare.restype = C.c_int
    lib.uwie_engine_check.argtypes = [C.c_void_p, C.c_void_p]
    lib.uwie_engine_check.restype = C.c_int
    lib.uwie_engine_last_launches.argtypes = [C.c_void_p]
    lib.uwie_engine_last_launches.restype = C.c_int
    lib.uwie_engine_graph_replays.argtypes = [C.c_void_p]
    lib.uwie_engine_graph_replays.restype = C.c_longlong
    lib.uwie_engine_debug.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_size_t)]
    lib.uwie_engine_debug.restype = C.c_void_p
    lib.uwie_engine_profile.argtypes = [C.c_void_p, C.c_int]
    lib.uwie_engine_profile.restype = C.c_int
    lib.uwie_engine_profile_read.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_double), C.POINTER(C.c_int),
                                             C.POINTER(C.c_int), C.c_int]
    lib.uwie_engine_profile_read.restype = C.c_int
    lib.uwie_vgg_workspace_bytes.argtypes = [C.c_int]
    lib.uwie_vgg_workspace_bytes.restype = C.c_size_t
    fp = C.POINTER(C.c_float)
    lib.uwie_vgg_enhance.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, fp, fp, fp, fp, C.c_void_p, C.c_void_p,
                                     C.c_size_t, C.c_void_p]
    lib.uwie_vgg_enhance.restype = C.c_int
    lib.uwie_vgg_enhance_dev.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                         C.c_void_p, C.c_size_t, C.c_void_p]
    lib.uwie_vgg_enhance_dev.restype = C.c_int
    lib.uwie_vgg_enhance_u8.argtypes = lib.uwie_vgg_enhance_dev.argtypes
    lib.uwie_vgg_enhance_u8.restype = C.c_int
    lib.uwie_vgg_enhance_bwd.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p,
                                         C.c_void_p]
    lib.uwie_vgg_enhance_bwd.restype = C.c_int
    lib.uwie_vgg_features_u8.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_size_t,
                                         C.c_void_p]
    lib.uwie_vgg_features_u8.restype = C.c_int
    for name in ("uwie_vgg_preprocess_u8", "uwie_vgg_preprocess_f32"):
        fn = getattr(lib, name)
        fn.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        fn.restype = C.c_int
    lib.uwie_features79_workspace_bytes.argtypes = [C.c_int, C.c_int, C.c_int]
    lib.uwie_features79_workspace_bytes.restype = C.c_size_t
    lib.uwie_features79.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
    lib.uwie_features79.restype = C.c_int
    lib.uwie_quality_workspace_bytes.argtypes = [C.c_int, C.c_int, C.c_int]
    lib.uwie_quality_workspace_bytes.restype = C.c_size_t
    lib.uwie_quality_scores.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double), C.c_void_p,
                                        C.c_void_p, C.c_size_t, C.c_void_p]
    lib.uwie_quality_scores.restype = C.c_int
    vp, ci, cd, cf, cz = C.c_void_p, C.c_int, C.c_double, C.c_float, C.c_size_t
    sigs = {
        "uwie_prim_transmission_head_u8": ([vp, ci, ci, ci, vp, cf, cf, ci, vp, vp], ci),
        "uwie_prim_clahe_lab_workspace_bytes": ([ci, ci, ci], cz),
        "uwie_prim_clahe_lab": ([vp, ci, ci, ci, cd, ci, ci, vp, vp, cz, vp], ci),
        "uwie_prim_equalize_hist": ([vp, ci, ci, ci, vp, vp, cz, vp], ci),
        "uwie_prim_equalize_hist_static": ([vp, ci, ci, ci, vp, vp], ci),
        "uwie_prim_quantise": ([vp, ci, cz, vp, vp], ci),
        "uwie_prim_atmospheric_light_workspace_bytes": ([ci, ci], cz),
        "uwie_prim_atmospheric_light": ([vp, ci, ci, ci, vp, vp, cz, vp], ci),
        "uwie_prim_guided_filter_workspace_bytes": ([ci, ci], cz),
        "uwie_prim_guided_filter": ([vp, vp, ci, ci, ci, cd, ci, vp, vp, cz, vp], ci),
        "uwie_prim_transmission_workspace_bytes": ([ci, ci], cz),
        "uwie_prim_transmission": ([vp, ci, ci, ci, vp, cd, ci, cd, cd, ci, vp, vp, cz, vp], ci),
        "uwie_prim_recover": ([vp, ci, ci, ci, vp, vp, vp, ci, vp], ci),
        "uwie_prim_percentile_stretch_workspace_bytes": ([ci], cz),
        "uwie_prim_percentile_stretch": ([vp, ci, ci, ci, cd, cd, cd, vp, vp, cz, vp], ci),
        "uwie_prim_gamma": ([vp, ci, cz, cd, ci, vp, vp], ci),
        "uwie_prim_scale_clip": ([vp, ci, ci, ci, C.POINTER(C.c_double), vp, vp], ci),
        "uwie_prim_channel_means": ([vp, ci, ci, ci, vp, vp, cz, vp], ci),
        "uwie_prim_region_score_workspace_bytes": ([ci, ci], cz),
        "uwie_prim_region_score": ([vp, ci, ci, ci, vp, vp, cz, vp], ci),
        "uwie_prim_brightest_pixel": ([vp, ci, ci, ci, vp, vp, cz, vp], ci),
    }
    for name, (args, res) in sigs.items():
        fn = getattr(lib, name)
        fn.argtypes, fn.restype = args, res
    lib.uwie_selftest_division.argtypes = [C.c_ulonglong, C.c_ulonglong, C.POINTER(C.c_ulonglong)]
    lib.uwie_selftest_division.restype = C.c_int
    _lib = lib
    return lib


def check(code: int) -> None:
    if code != 0:
        raise UwieError(code, load().uwie_last_error().decode("utf-8", "replace"))


def preset(library_id: int):
    lib = Library()
    arr = (Strategy * MAX_STRATEGIES)()
    n = load().uwie_preset(library_id, C.byref(lib), arr)
    if n < 0:
        check(n)
    return lib, [arr[i] for i in range(n)]


# ---------------------------------------------------------------------------------------------
# Scratch memory for the entry points.  Two test hooks replace the tools this GPU pool keeps closed (compute-sanitizer):
#   UWIE_POISON=<0..255>   every scratch buffer is filled with that byte before use: results must not depend on
#                          uninitialised scratch (initcheck's job);
#   UWIE_GUARD=<bytes>     every scratch buffer gets that many canary bytes before and after it; guards_intact() tells
#                          whether any kernel wrote outside what the *_workspace_bytes functions promised (memcheck's job).
# Both are off in normal use and cost nothing then.
# ---------------------------------------------------------------------------------------------
_CANARY = 0xC3
_GUARDED = []


def scratch(nbytes: int, device):
    import torch
    nbytes = max(int(nbytes), 256)
    guard = int(os.environ.get("UWIE_GUARD", "0") or 0)
    guard = (guard + 255) & ~255
    poison = os.environ.get("UWIE_POISON", "")
    if not guard and poison == "":
        return torch.empty(nbytes, dtype=torch.uint8, device=device)
    padded = (nbytes + 255) & ~255
    whole = torch.empty(padded + 2 * guard, dtype=torch.uint8, device=device)
    if guard:
        whole[:guard] = _CANARY
        whole[guard + padded:] = _CANARY
        _GUARDED.append((whole, guard, padded))
    inner = whole[guard:guard + nbytes]
    if poison != "":
        whole[guard:guard + padded] = int(poison) & 255
    return inner


def guards_intact(clear: bool = True) -> bool:
    """True when no canary byte around any scratch buffer handed out since the last call was overwritten."""
    import torch
    ok = True
    for whole, guard, padded in _GUARDED:
        torch.cuda.synchronize(whole.device)
        ok = ok and bool((whole[:guard] == _CANARY).all()) and bool((whole[guard + padded:] == _CANARY).all())
    if clear:
        _GUARDED.clear()
    return ok
