"""Memory-mapped GGUF v3 reader.

Capability parity: `llama.cpp/ggml/src/gguf.cpp` (parse header / KV / tensor infos / aligned blob) and
split files `%s-%05d-of-%05d.gguf` (`llama.cpp/src/llama.cpp:1094-1100`).  Design differs on purpose:
the file is mmapped once and every tensor is a zero-copy numpy view, so each rank can slice *its own
shard* straight out of the page cache and upload it (no weight traffic over a socket).
"""
from __future__ import annotations

import mmap
import os
import re
import struct
from dataclasses import dataclass
from typing import Any, Iterator

import numpy as np

from .constants import (BLOCK_GEOMETRY, GGUF_DEFAULT_ALIGNMENT, GGUF_MAGIC, GGMLType, SCALAR_DTYPES, ValueType,
                        row_bytes)
from .quants import dequantize


@dataclass
class TensorInfo:
    name: str
    shape: tuple[int, ...]  # logical row-major shape (outermost first) - GGUF stores ne[] innermost first
    ggml_type: GGMLType
    offset: int  # relative to data section
    nbytes: int
    file_index: int = 0

    @property
    def n_elements(self) -> int:
        n = 1
        for s in self.shape:
            n *= s
        return n

    @property
    def row_nbytes(self) -> int:
        return row_bytes(self.ggml_type, self.shape[-1])


class _Cursor:
    def __init__(self, buf, pos: int = 0):
        self.buf = buf
        self.pos = pos

    def take(self, fmt: str):
        size = struct.calcsize(fmt)
        if self.pos + size > len(self.buf):
            raise ValueError("GGUF: truncated file")
        (v,) = struct.unpack_from(fmt, self.buf, self.pos)
        self.pos += size
        return v

    def string(self) -> str:
        n = self.take("<Q")
        if self.pos + n > len(self.buf):
            raise ValueError("GGUF: truncated string")
        s = bytes(self.buf[self.pos:self.pos + n]).decode("utf-8", errors="replace")
        self.pos += n
        return s

    def value(self, vtype: ValueType) -> Any:
        if vtype == ValueType.STRING:
            return self.string()
        if vtype == ValueType.ARRAY:
            etype = ValueType(self.take("<I"))
            count = self.take("<Q")
            if etype == ValueType.STRING:
                return [self.string() for _ in range(count)]
            if etype == ValueType.ARRAY:
                return [self.value(ValueType.ARRAY) for _ in range(count)]
            dt = SCALAR_DTYPES[etype]
            nbytes = dt.itemsize * count
            arr = np.frombuffer(self.buf, dtype=dt, count=count, offset=self.pos).copy()
            self.pos += nbytes
            return arr
        dt = SCALAR_DTYPES[vtype]
        v = np.frombuffer(self.buf, dtype=dt, count=1, offset=self.pos)[0]
        self.pos += dt.itemsize
        return v.item()


class GGUFFile:
    """One GGUF file (or a set of split shards opened through :func:`open_gguf`)."""

    def __init__(self, path: str, use_mmap: bool | None = None, mlock: bool | None = None, direct_io: bool | None = None):
        """use_mmap=False reads the file into anonymous memory instead of mapping it (`--no-mmap`), direct_io=True does that with
        O_DIRECT so the page cache is bypassed (`--direct-io`), mlock=True pins the pages (`--mlock`) - the loader options of
        `llama.cpp/src/llama-mmap.cpp` (llama_mmap / llama_mlock / llama_file direct I/O).  None: DLP_NO_MMAP / DLP_DIRECT_IO /
        DLP_MLOCK from the environment (how the command-line flags reach every tool)."""
        self.path = path
        use_mmap = os.environ.get("DLP_NO_MMAP", "0") != "1" if use_mmap is None else use_mmap
        direct_io = os.environ.get("DLP_DIRECT_IO", "0") == "1" if direct_io is None else direct_io
        mlock = os.environ.get("DLP_MLOCK", "0") == "1" if mlock is None else mlock
        self._fh = open(path, "rb")
        size = os.fstat(self._fh.fileno()).st_size
        if size < 24:
            raise ValueError(f"{path}: not a GGUF file (too small)")
        self.load_mode = "mmap"
        if use_mmap and not direct_io:
            self._mm = mmap.mmap(self._fh.fileno(), 0, access=mmap.ACCESS_READ)
        else:
            self._mm = self._read_all(path, size, direct_io)
        self.mlocked = self._mlock(size) if mlock else False
        cur = _Cursor(self._mm)
        if cur.take("<I") != GGUF_MAGIC:
            raise ValueError(f"{path}: bad magic")
        self.version = cur.take("<I")
        if self.version not in (2, 3):
            raise ValueError(f"{path}: unsupported GGUF version {self.version}")
        n_tensors = cur.take("<Q")
        n_kv = cur.take("<Q")
        self.metadata: dict[str, Any] = {}
        self.metadata_types: dict[str, ValueType] = {}
        for _ in range(n_kv):
            key = cur.string()
            vtype = ValueType(cur.take("<I"))
            self.metadata[key] = cur.value(vtype)
            self.metadata_types[key] = vtype
        # `--override-kv KEY=TYPE:VALUE` (and the rope / YaRN flags built on it): replaces metadata values of model files for every loader in
        # the process; only files that describe a model (they carry an architecture) and not adapters / importance matrices are touched
        if os.environ.get("DLP_OVERRIDE_KV") and "general.architecture" in self.metadata and self.metadata.get("general.type", "model") == "model":
            from ..utils.args import parse_override_kv
            for spec in os.environ["DLP_OVERRIDE_KV"].split(";"):
                if spec.strip():
                    k, v = parse_override_kv(spec, str(self.metadata["general.architecture"]))
                    self.metadata[k] = v
        self.alignment = int(self.metadata.get("general.alignment", GGUF_DEFAULT_ALIGNMENT))
        self.tensors: dict[str, TensorInfo] = {}
        for _ in range(n_tensors):
            name = cur.string()
            n_dims = cur.take("<I")
            ne = [cur.take("<Q") for _ in range(n_dims)]
            ttype = GGMLType(cur.take("<I"))
            offset = cur.take("<Q")
            shape = tuple(reversed(ne)) if ne else (1,)
            per, nb = BLOCK_GEOMETRY[ttype]
            n_el = 1
            for s in shape:
                n_el *= s
            if shape[-1] % per:
                raise ValueError(f"{path}: tensor {name} row length {shape[-1]} not a multiple of {per}")
            self.tensors[name] = TensorInfo(name, shape, ttype, offset, n_el // per * nb)
        self.data_offset = (cur.pos + self.alignment - 1) // self.alignment * self.alignment
        for t in self.tensors.values():
            if self.data_offset + t.offset + t.nbytes > size:
                raise ValueError(f"{path}: tensor {t.name} extends past end of file")

    def _read_all(self, path: str, size: int, direct_io: bool):
        """The whole file in page-aligned anonymous memory; with direct_io the reads bypass the page cache (block-aligned chunks
        through O_DIRECT, the unaligned tail through the ordinary descriptor)."""
        buf = mmap.mmap(-1, max(size, 1))
        view = memoryview(buf)
        done = 0
        self.load_mode = "read"
        if direct_io and hasattr(os, "O_DIRECT"):
            try:
                fd = os.open(path, os.O_RDONLY | os.O_DIRECT)
            except OSError:
                fd = -1          # tmpfs and friends refuse O_DIRECT: fall back to buffered reads
            if fd >= 0:
                try:
                    blk, chunk = 4096, 64 << 20
                    aligned = size // blk * blk
                    while done < aligned:
                        n = os.preadv(fd, [view[done:min(aligned, done + chunk)]], done)
                        if n <= 0:
                            break
                        done += n
                    if done == aligned:
                        self.load_mode = "direct"
                    else:
                        done = 0
                except OSError:
                    done = 0
                finally:
                    os.close(fd)
        self._fh.seek(done)
        while done < size:
            n = self._fh.readinto(view[done:])
            if not n:
                raise ValueError(f"{path}: short read")
            done += n
        return buf

    def _mlock(self, size: int) -> bool:
        import ctypes
        try:
            addr = np.frombuffer(self._mm, dtype=np.uint8, count=1).ctypes.data
            libc = ctypes.CDLL(None, use_errno=True)
            if libc.mlock(ctypes.c_void_p(addr), ctypes.c_size_t(size)) != 0:
                err = ctypes.get_errno()
                import warnings
                warnings.warn(f"mlock of {size / 2**20:.0f} MiB failed ({os.strerror(err)}): raise RLIMIT_MEMLOCK (ulimit -l) - continuing unlocked")
                return False
            return True
        except (OSError, AttributeError, ValueError):
            return False

    # ---- tensor access --------------------------------------------------------------------
    def raw(self, name: str) -> np.ndarray:
        """Zero-copy uint8 view of the tensor's bytes, shaped [rows..., row_nbytes]."""
        t = self.tensors[name]
        flat = np.frombuffer(self._mm, dtype=np.uint8, count=t.nbytes, offset=self.data_offset + t.offset)
        return flat.reshape(t.shape[:-1] + (t.row_nbytes,))

    def rows_raw(self, name: str, start: int, stop: int) -> np.ndarray:
        """Raw bytes of outermost-dim rows [start, stop) of a 2-D (or stacked) tensor."""
        return self.raw(name)[start:stop]

    def array(self, name: str, dtype=np.float32) -> np.ndarray:
        """Decoded tensor as a dense array (float32 by default)."""
        t = self.tensors[name]
        if t.ggml_type == GGMLType.F32 and dtype == np.float32:
            return np.frombuffer(self._mm, "<f4", t.n_elements, self.data_offset + t.offset).reshape(t.shape)
        if t.ggml_type == GGMLType.F16 and dtype == np.float16:
            return np.frombuffer(self._mm, "<f2", t.n_elements, self.data_offset + t.offset).reshape(t.shape)
        return dequantize(self.raw(name), t.ggml_type, t.shape).astype(dtype, copy=False)

    def close(self):
        try:
            self._mm.close()
        except BufferError:
            pass  # live numpy views keep the mapping alive; the OS reclaims it with the process
        self._fh.close()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


_SPLIT_RE = re.compile(r"^(.*)-(\d{5})-of-(\d{5})\.gguf$")


class GGUFModelFiles:
    """A model that may span several split shards; exposes the union of tensors and shard-0 metadata."""

    def __init__(self, files: list[GGUFFile]):
        self.files = files
        self.metadata = files[0].metadata
        self.tensors: dict[str, TensorInfo] = {}
        for i, f in enumerate(files):
            for name, t in f.tensors.items():
                t.file_index = i
                self.tensors[name] = t

    def _file(self, name: str) -> GGUFFile:
        return self.files[self.tensors[name].file_index]

    def raw(self, name: str) -> np.ndarray:
        return self._file(name).raw(name)

    def rows_raw(self, name: str, start: int, stop: int) -> np.ndarray:
        return self._file(name).rows_raw(name, start, stop)

    def array(self, name: str, dtype=np.float32) -> np.ndarray:
        return self._file(name).array(name, dtype)

    def __contains__(self, name: str) -> bool:
        return name in self.tensors

    def names(self) -> Iterator[str]:
        return iter(self.tensors)

    def close(self):
        for f in self.files:
            f.close()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def open_gguf(path: str) -> GGUFModelFiles:
    """Open a GGUF checkpoint; if `path` is shard 1 of a `-00001-of-0000N.gguf` set, open all shards."""
    m = _SPLIT_RE.match(path)
    if not m:
        return GGUFModelFiles([GGUFFile(path)])
    prefix, _, total = m.group(1), int(m.group(2)), int(m.group(3))
    files = [GGUFFile(f"{prefix}-{i:05d}-of-{total:05d}.gguf") for i in range(1, total + 1)]
    return GGUFModelFiles(files)
