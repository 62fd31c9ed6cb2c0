"""Build + locate the native (C++) runtime binaries: dlp-orchestrator, dlp-worker."""
from __future__ import annotations

import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
BIN = os.path.join(HERE, "bin")
TARGETS = {
    "dlp-orchestrator": (["orchestrator.cpp"], ["-pthread"]),
    "dlp-worker": (["worker.cpp"], ["-pthread", "-fopenmp", "-mavx2", "-mfma", "-mf16c"]),
}


# Sanitizer builds (parity: the reference's LLAMA_SANITIZE_{ADDRESS,THREAD,UNDEFINED} CMake switches): the same
# sources compiled with -fsanitize=... into bin/<kind>/; tests/test_sanitizers.py drives them with real traffic.
SANITIZERS = {"address": ["-fsanitize=address,undefined", "-fno-omit-frame-pointer", "-g", "-O1"],
              "thread": ["-fsanitize=thread", "-fno-omit-frame-pointer", "-g", "-O1"]}


# dlp-run links the CUDA runtime and the in-tree kernel library (found next to it through an rpath): built when both exist
CUDA_HOME = os.environ.get("CUDA_HOME", "/usr/local/cuda")
OPS_DIR = os.path.normpath(os.path.join(HERE, "..", "ops"))


def _refresh_targets():
    """(evaluated at build time, not import time: the kernel library may be built after this module was imported)"""
    if os.path.exists(os.path.join(CUDA_HOME, "include", "cuda_runtime.h")) and os.path.exists(os.path.join(OPS_DIR, "libdlp_b200.so")):
        TARGETS["dlp-run"] = (["engine.cpp"], ["-pthread", "-mavx2", "-mf16c", f"-I{CUDA_HOME}/include", f"-L{CUDA_HOME}/lib64", f"-L{OPS_DIR}",
                                               "-Wl,-rpath," + OPS_DIR, "-Wl,-rpath," + f"{CUDA_HOME}/lib64", "-ldlp_b200", "-lcudart"])


def path(name: str, sanitize: str | None = None) -> str:
    return os.path.join(BIN, sanitize, name) if sanitize else os.path.join(BIN, name)


def build(force: bool = False, sanitize: str | None = None) -> dict[str, str]:
    if sanitize and sanitize not in SANITIZERS:
        raise ValueError(f"unknown sanitizer {sanitize!r}; known: {sorted(SANITIZERS)}")
    os.makedirs(os.path.join(BIN, sanitize) if sanitize else BIN, exist_ok=True)
    _refresh_targets()
    out = {}
    for name, (srcs, flags) in TARGETS.items():
        src_paths = [os.path.join(HERE, s) for s in srcs]
        if not all(os.path.exists(s) for s in src_paths):
            continue
        dst = path(name, sanitize)
        deps = src_paths + [os.path.join(HERE, f) for f in os.listdir(HERE) if f.endswith(".h")]
        if not force and os.path.exists(dst) and all(os.path.getmtime(dst) >= os.path.getmtime(d) for d in deps):
            out[name] = dst
            continue
        opt = SANITIZERS[sanitize] if sanitize else ["-O3"]
        libs = [f for f in flags if f.startswith("-l")]
        cmd = ["g++", "-std=c++17", "-Wall", "-Wno-unused-function"] + opt + [f for f in flags if not f.startswith("-l")] + src_paths + ["-o", dst] + libs
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"building {name} failed:\n{res.stderr}")
        out[name] = dst
    return out
