"""Makes the reference arm available: the UNMODIFIED llama.cpp of /root/reference, built with its own CMake for sm_100.

`baseline/_ref/bin` (git-ignored) holds `llama-bench`, `llama-batched-bench`, `llama-cli`, `llama-server`, `rpc-server`,
`llama-perplexity` and their shared libraries.  Three ways to get it, tried in this order:

1. it is already there;
2. `baseline/ref_prebuilt.tar.xz` (tracked): the output of `tools/build_reference.sh` on this image, archived because a
   fresh checkout on a GPU box has neither `/root/reference` nor the ten minutes of 8-core nvcc time the build needs -
   without it the driver's reference arm reports "unavailable" and the headline ratio cannot be formed;
3. `tools/build_reference.sh` from `/root/reference/llama.cpp` when that tree and cmake exist (optional time limit).

Nothing of this repository is linked into or loaded by those binaries.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import tarfile

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
BIN = os.path.join(HERE, "_ref", "bin")
ARCHIVE = os.path.join(HERE, "ref_prebuilt.tar.xz")
TOOLS = ("llama-bench", "llama-batched-bench", "llama-cli", "llama-server", "rpc-server", "llama-perplexity")


def have(tool: str = "llama-bench") -> bool:
    return os.path.exists(os.path.join(BIN, tool))


def exe(tool: str) -> str:
    return os.path.join(BIN, tool)


def env() -> dict:
    e = dict(os.environ)
    e["LD_LIBRARY_PATH"] = BIN + ":" + e.get("LD_LIBRARY_PATH", "")
    return e


def cpu_env(scratch: str) -> dict:
    """Environment for running the reference's ggml-cpu back-end (`-ngl 0`) on a machine without an NVIDIA driver.

    The binaries are the CUDA build, so the loader wants `libcuda.so.1`; where the driver library is absent the toolkit's link
    stub stands in (its `cuInit` fails, ggml reports "no CUDA devices" and every tensor stays on the CPU back-end).
    """
    import ctypes
    e = env()
    try:
        ctypes.CDLL("libcuda.so.1")
    except OSError:
        stub = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "lib64", "stubs", "libcuda.so")
        if os.path.exists(stub):
            os.makedirs(scratch, exist_ok=True)
            link = os.path.join(scratch, "libcuda.so.1")
            if not os.path.exists(link):
                os.symlink(stub, link)
            e["LD_LIBRARY_PATH"] = e["LD_LIBRARY_PATH"].rstrip(":") + ":" + scratch
    e["CUDA_VISIBLE_DEVICES"] = ""
    return e


def ensure(allow_build: bool = True, build_timeout_s: float | None = None, log=lambda s: None) -> str | None:
    """Returns None when the reference is in place, else a one-line reason."""
    if have():
        return None
    if os.path.exists(ARCHIVE):
        try:
            os.makedirs(os.path.dirname(BIN), exist_ok=True)
            with tarfile.open(ARCHIVE, "r:xz") as t:
                t.extractall(os.path.dirname(BIN))
            if have():
                log("reference: unpacked baseline/ref_prebuilt.tar.xz")
                return None
        except (OSError, tarfile.TarError) as e:
            log(f"reference: archive unusable ({e})")
    src = os.environ.get("DLP_REFERENCE_SRC", "/root/reference/llama.cpp")
    if not allow_build:
        return "baseline/_ref/bin/llama-bench absent and building was not allowed"
    if not os.path.isdir(src):
        return f"baseline/_ref/bin/llama-bench absent, no archive, and no reference source at {src}"
    if shutil.which("cmake") is None:
        return "baseline/_ref/bin/llama-bench absent and cmake is not installed"
    try:
        p = subprocess.run(["bash", os.path.join(REPO, "tools", "build_reference.sh")], capture_output=True, text=True,
                           timeout=build_timeout_s)
    except subprocess.TimeoutExpired:
        return f"reference build exceeded {build_timeout_s:.0f} s"
    if p.returncode != 0 or not have():
        return f"reference build failed: {(p.stderr or p.stdout).strip()[-200:]}"
    return None


def pack() -> str:
    """Archive baseline/_ref/bin into the tracked tarball (run after tools/build_reference.sh)."""
    assert have(), "build the reference first"
    with tarfile.open(ARCHIVE, "w:xz", preset=6) as t:
        t.add(BIN, arcname="bin")
    return ARCHIVE


if __name__ == "__main__":
    import sys
    if len(sys.argv) > 1 and sys.argv[1] == "pack":
        print(pack())
    else:
        why = ensure(log=print)
        print("ok" if why is None else why)
        sys.exit(0 if why is None else 1)
