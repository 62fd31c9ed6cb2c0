"""ctypes binding of libdlp_b200.so (C ABI in include/dlp_b200.h).

The library is built in-tree by `csrc/build.sh` (nvcc, sm_100a) so it travels with the repository snapshot
and shows up as a loaded in-tree `.so`.  On a machine with a GPU the native path is mandatory: a missing
library raises instead of silently falling back.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
_REPO = os.path.dirname(os.path.dirname(_HERE))
LIB_PATH = os.path.join(_HERE, "libdlp_b200.so")

_lib = None
_lock = threading.Lock()


class dlp_weight(C.Structure):
    _fields_ = [("base", C.c_void_p), ("aux0", C.c_void_p), ("aux1", C.c_void_p), ("aux2", C.c_void_p),
                ("wtype", C.c_int32), ("N", C.c_int32), ("K", C.c_int32), ("reserved", C.c_int32)]


def build(force: bool = False) -> str:
    """Compile the CUDA sources for sm_100a (cross-compiles without a GPU)."""
    script = os.path.join(_REPO, "csrc", "build.sh")
    if force and os.path.isdir(os.path.join(_REPO, "csrc", "build")):
        for f in os.listdir(os.path.join(_REPO, "csrc", "build")):
            if f.endswith(".o"):
                os.remove(os.path.join(_REPO, "csrc", "build", f))
    res = subprocess.run(["bash", script], capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"building libdlp_b200.so failed:\n{res.stdout}\n{res.stderr}")
    return LIB_PATH


def _declare(lib):
    vp, i32, f32, u32, sz = C.c_void_p, C.c_int, C.c_float, C.c_uint32, C.c_size_t
    W = C.POINTER(dlp_weight)
    sigs = {
        "dlp_last_error": (C.c_char_p, []),
        "dlp_version": (i32, []),
        "dlp_sm_count": (i32, []),
        "dlp_gemv": (i32, [W, W, vp, vp, f32, vp, i32, i32, i32, vp]),
        "dlp_gemm_f16": (i32, [vp, vp, i32, vp, i32, vp, i32, i32, i32, i32, i32, i32, i32, vp]),
        "dlp_linear": (i32, [W, W, vp, vp, f32, vp, i32, i32, i32, vp, sz, vp]),
        "dlp_dequant_rows_f32": (i32, [W, vp, i32, vp, i32, vp]),
        "dlp_dequant_rows_f16": (i32, [W, vp, i32, vp, i32, vp]),
        "dlp_rmsnorm": (i32, [vp, vp, vp, vp, i32, i32, f32, vp]),
        "dlp_layernorm": (i32, [vp, vp, vp, vp, vp, i32, i32, f32, vp]),
        "dlp_bias_act": (i32, [vp, vp, vp, i32, i32, i32, vp]),
        "dlp_rope_kv": (i32, [vp, i32, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, f32, vp]),
        "dlp_rope_kv_ex": (i32, [vp, i32, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, f32, vp, vp, vp, f32,
                                 vp, C.c_size_t, vp, C.c_size_t, vp]),
        "dlp_swiglu": (i32, [vp, i32, vp, i32, i32, i32, vp]),
        "dlp_advance": (i32, [vp, vp, vp, vp, vp, vp, i32, i32, i32, vp]),
        "dlp_attn_paged": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp]),
        "dlp_attn_paged_fused": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp]),
        "dlp_set_trace": (i32, [vp]),
        "dlp_advance_rope": (i32, [vp, vp, vp, vp, vp, vp, i32, i32, i32, vp, i32, vp, vp]),
        "dlp_attn_decode_rope": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, f32, vp]),
        "dlp_attn_prefill": (i32, [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp]),
        "dlp_argmax": (i32, [vp, i32, i32, i32, vp, vp]),
        "dlp_sample": (i32, [vp, i32, i32, i32, i32, f32, f32, f32, u32, vp, vp, vp]),
        "dlp_f32_to_f16": (i32, [vp, vp, sz, vp]),
        "dlp_moe_route": (i32, [vp, i32, i32, i32, vp, vp, vp]),
        "dlp_gemv_moe": (i32, [W, W, vp, vp, f32, vp, i32, i32, vp, i32, vp, vp]),
    }
    for name, (res, args) in sigs.items():
        fn = getattr(lib, name, None)
        if fn is None:
            continue  # optional / newer entry points are declared where they are used
        fn.restype = res
        fn.argtypes = args


def available() -> bool:
    return os.path.exists(LIB_PATH)


def load():
    """Return the loaded library, building it first if the `.so` is absent and nvcc exists."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                build()
            lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
            _declare(lib)
            _lib = lib
    return _lib


class DlpError(RuntimeError):
    pass


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = load().dlp_last_error()
        raise DlpError(f"{what} failed (rc={rc}): {msg.decode() if msg else ''}")
