"""ctypes binding of include/pocket_tts_b200.h.  The product path has no fallback: if the
shared library is missing or fails to load, importing this module raises."""

from __future__ import annotations

import ctypes as C
import os
import re
from pathlib import Path

HERE = Path(__file__).resolve().parent
# PTTS_NATIVE_LIB: developer override, an alternative build of the same library (A/B timing of compile-time choices)
LIB_PATH = Path(os.environ["PTTS_NATIVE_LIB"]) if os.environ.get("PTTS_NATIVE_LIB") else HERE / "libpocket_tts_b200.so"
HEADER = HERE.parent / "include" / "pocket_tts_b200.h"

PTTS_OK, PTTS_EINVAL, PTTS_ECUDA, PTTS_ESTATE, PTTS_ENOSLOT = 0, -1, -2, -3, -4
PRECISION_FP32, PRECISION_BF16 = 0, 1


class PttsConfig(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "n_layers", "d_model", "n_heads", "ffn", "ldim", "flow_dim", "flow_depth", "mimi_d",
        "mimi_heads", "mimi_layers", "mimi_ffn", "mimi_context", "n_filters")] + \
        [("ratios", C.c_int32 * 3)] + \
        [(n, C.c_int32) for n in (
            "kernel", "res_kernel", "last_kernel", "upsample_stride", "max_streams", "kv_capacity",
            "max_prefill_rows", "max_lsd_steps", "precision", "device", "kv_fp8")]


class NativeError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"pocket_tts_b200 native error {code}: {msg}")
        self.code = code


def exported_symbols_from_header() -> list[str]:
    """Every function the header declares (used by the CPU test that checks the .so)."""
    text = HEADER.read_text()
    return sorted(set(re.findall(r"\b(ptts_[a-z_0-9]+)\s*\(", text)))


def load_library() -> C.CDLL:
    if not LIB_PATH.exists():
        raise ImportError(
            f"{LIB_PATH} not found: build it with `python -m pocket_tts_b200.build` "
            "(there is no CPU / PyTorch fallback for the generation path)")
    lib = C.CDLL(str(LIB_PATH))
    p, i32, i64, f32p, i32p = C.c_void_p, C.c_int32, C.c_int64, C.c_void_p, C.POINTER(C.c_int32)
    sig = {
        "ptts_last_error": (C.c_char_p, []),
        "ptts_version": (C.c_int, []),
        "ptts_engine_create": (C.c_int, [C.POINTER(PttsConfig), C.POINTER(p)]),
        "ptts_engine_destroy": (C.c_int, [p]),
        "ptts_load_tensor": (C.c_int, [p, C.c_char_p, f32p, C.POINTER(i64), i32]),
        "ptts_finalize": (C.c_int, [p]),
        "ptts_stream_open": (C.c_int, [p, i32p]),
        "ptts_stream_close": (C.c_int, [p, i32]),
        "ptts_stream_reset_decoder": (C.c_int, [p, i32]),
        "ptts_stream_fork": (C.c_int, [p, i32, i32]),
        "ptts_stream_fork_shared": (C.c_int, [p, i32, i32]),
        "ptts_stream_import_kv": (C.c_int, [p, i32, i32, f32p, i32]),
        "ptts_stream_export_kv": (C.c_int, [p, i32, i32, f32p, i32]),
        "ptts_stream_set_offset": (C.c_int, [p, i32, i32]),
        "ptts_stream_get_offset": (C.c_int, [p, i32, i32p]),
        "ptts_stream_set_prev_latent": (C.c_int, [p, i32, f32p]),
        "ptts_prefill": (C.c_int, [p, i32, i32p, i32p, f32p]),
        "ptts_prefill_tokens": (C.c_int, [p, i32, i32p, i32p, i32p]),
        "ptts_step": (C.c_int, [p, i32, i32p, f32p, i32, i32, f32p, f32p, f32p]),
        "ptts_mimi_decode": (C.c_int, [p, i32, i32p, f32p, f32p]),
        "ptts_mimi_decode_frames": (C.c_int, [p, i32, i32p, i32, f32p, f32p]),
        "ptts_run_frames": (C.c_int, [p, i32, i32p, i32, f32p, i32, f32p, f32p]),
        "ptts_generate_frames": (C.c_int, [p, i32, i32p, i32, f32p, i32, f32p, f32p, f32p]),
        "ptts_encode_audio": (C.c_int, [p, f32p, i64, f32p, i32, i32p]),
        "ptts_pcm_to_int16": (C.c_int, [p, f32p, p, i64]),
        "ptts_pcm_to_int16_static": (C.c_int, [p, f32p, p, i64]),
        "ptts_synchronize": (C.c_int, [p]),
        "ptts_launch_count": (i64, [p]),
        "ptts_stream_handle": (p, [p]),
        "ptts_set_graphs": (C.c_int, [p, i32]),
        "ptts_debug_read": (C.c_int, [p, C.c_char_p, f32p, i64, C.POINTER(i64)]),
        "ptts_debug_gemm": (C.c_int, [p, i32, i32, i32, i32, i32, i32, i32, f32p, i64, f32p, f32p]),
        "ptts_profile_begin": (C.c_int, [p]),
        "ptts_profile_end": (C.c_int, [p, C.c_char_p, i64]),
        "ptts_debug_timeline": (C.c_int, [p, C.c_char_p, i64, i32]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return lib


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        _lib = load_library()
    return _lib


def check(rc: int) -> None:
    if rc != PTTS_OK:
        msg = lib().ptts_last_error()
        raise NativeError(rc, msg.decode("utf-8", "replace") if msg else "")
