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


def path(name: str) -> str:
    return os.path.join(BIN, name)


def build(force: bool = False) -> dict[str, str]:
    os.makedirs(BIN, exist_ok=True)
    out = {}
    for name, (srcs, flags) in TARGETS.items():
        src_paths = [os.path.join(HERE, s) for s in srcs]
        if not all(os.path.exists(s) for s in src_paths):
            continue
        dst = path(name)
        deps = src_paths + [os.path.join(HERE, f) for f in os.listdir(HERE) if f.endswith(".h")]
        if not force and os.path.exists(dst) and all(os.path.getmtime(dst) >= os.path.getmtime(d) for d in deps):
            out[name] = dst
            continue
        cmd = ["g++", "-O3", "-std=c++17", "-Wall", "-Wno-unused-function"] + flags + src_paths + ["-o", dst]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"building {name} failed:\n{res.stderr}")
        out[name] = dst
    return out
