"""Streaming GGUF v3 writer.

Capability parity: `gguf_write_to_file` (`llama.cpp/ggml/src/gguf.cpp`) / gguf-py `GGUFWriter`.
Design: the layout (offsets) is computed from declared shapes first; tensor payloads are then produced
one at a time by callbacks so that a 16 GB random-init checkpoint never has to exist in host memory.
"""
from __future__ import annotations

import struct
from typing import Any, Callable

import numpy as np

from .constants import BLOCK_GEOMETRY, GGUF_DEFAULT_ALIGNMENT, GGUF_MAGIC, GGUF_VERSION, GGMLType, ValueType

_PACK = {
    ValueType.UINT8: "<B", ValueType.INT8: "<b", ValueType.UINT16: "<H", ValueType.INT16: "<h",
    ValueType.UINT32: "<I", ValueType.INT32: "<i", ValueType.FLOAT32: "<f", ValueType.BOOL: "<?",
    ValueType.UINT64: "<Q", ValueType.INT64: "<q", ValueType.FLOAT64: "<d",
}
_NP = {
    ValueType.UINT8: "<u1", ValueType.INT8: "<i1", ValueType.UINT16: "<u2", ValueType.INT16: "<i2",
    ValueType.UINT32: "<u4", ValueType.INT32: "<i4", ValueType.FLOAT32: "<f4", ValueType.BOOL: "?",
    ValueType.UINT64: "<u8", ValueType.INT64: "<i8", ValueType.FLOAT64: "<f8",
}


def _s(text: str) -> bytes:
    b = text.encode("utf-8")
    return struct.pack("<Q", len(b)) + b


def _infer_type(v: Any) -> ValueType:
    if isinstance(v, bool):
        return ValueType.BOOL
    if isinstance(v, int):
        return ValueType.UINT32 if 0 <= v < 2 ** 32 else ValueType.INT64
    if isinstance(v, float):
        return ValueType.FLOAT32
    if isinstance(v, str):
        return ValueType.STRING
    raise TypeError(f"cannot infer GGUF type for {type(v)}")


class GGUFWriter:
    def __init__(self, path: str, alignment: int = GGUF_DEFAULT_ALIGNMENT):
        self.path = path
        self.alignment = alignment
        self._kv: list[tuple[str, ValueType, Any, ValueType | None]] = []
        self._tensors: list[tuple[str, tuple[int, ...], GGMLType, Callable[[], np.ndarray]]] = []

    # ---- metadata -------------------------------------------------------------------------
    def add(self, key: str, value: Any, vtype: ValueType | None = None, etype: ValueType | None = None):
        if isinstance(value, (list, tuple, np.ndarray)):
            if etype is None:
                if isinstance(value, np.ndarray):
                    etype = next(vt for vt, s in _NP.items() if np.dtype(s) == value.dtype)
                else:
                    etype = _infer_type(value[0]) if len(value) else ValueType.UINT32
            self._kv.append((key, ValueType.ARRAY, value, etype))
        else:
            self._kv.append((key, vtype or _infer_type(value), value, None))

    def _encode_kv(self) -> bytes:
        out = bytearray()
        for key, vtype, value, etype in self._kv:
            out += _s(key) + struct.pack("<I", int(vtype))
            if vtype == ValueType.STRING:
                out += _s(value)
            elif vtype == ValueType.ARRAY:
                out += struct.pack("<IQ", int(etype), len(value))
                if etype == ValueType.STRING:
                    out += b"".join(_s(x) for x in value)
                else:
                    out += np.asarray(value).astype(_NP[etype]).tobytes()
            else:
                out += struct.pack(_PACK[vtype], value)
        return bytes(out)

    # ---- tensors --------------------------------------------------------------------------
    def add_tensor(self, name: str, shape: tuple[int, ...], ggml_type: GGMLType, producer: Callable[[], np.ndarray]):
        """`shape` is row-major (outermost first); `producer()` returns the raw payload (uint8-viewable)."""
        per, _ = BLOCK_GEOMETRY[GGMLType(ggml_type)]
        if shape[-1] % per:
            raise ValueError(f"{name}: row length {shape[-1]} not a multiple of block size {per}")
        self._tensors.append((name, tuple(int(s) for s in shape), GGMLType(ggml_type), producer))

    def add_array(self, name: str, arr: np.ndarray, ggml_type: GGMLType | None = None):
        """Convenience: store a dense numpy array (F32/F16) or quantise it to `ggml_type`."""
        from .quants import quantize

        if ggml_type is None:
            ggml_type = {np.dtype(np.float32): GGMLType.F32, np.dtype(np.float16): GGMLType.F16}[arr.dtype]
        shape = arr.shape
        self.add_tensor(name, shape, ggml_type, lambda a=arr, t=ggml_type: quantize(np.asarray(a, np.float32), t))

    @staticmethod
    def _nbytes(shape, ggml_type) -> int:
        per, nb = BLOCK_GEOMETRY[ggml_type]
        n = 1
        for s in shape:
            n *= s
        return n // per * nb

    def _pad(self, n: int) -> int:
        return (n + self.alignment - 1) // self.alignment * self.alignment

    def write(self):
        if self.alignment != GGUF_DEFAULT_ALIGNMENT and not any(k == "general.alignment" for k, *_ in self._kv):
            self.add("general.alignment", self.alignment, ValueType.UINT32)
        kv = self._encode_kv()
        infos = bytearray()
        offset = 0
        sizes = []
        for name, shape, ttype, _ in self._tensors:
            nbytes = self._nbytes(shape, ttype)
            infos += _s(name) + struct.pack("<I", len(shape))
            for dim in reversed(shape):  # GGUF stores ne[] innermost first
                infos += struct.pack("<Q", dim)
            infos += struct.pack("<IQ", int(ttype), offset)
            sizes.append(nbytes)
            offset = self._pad(offset + nbytes)
        header = struct.pack("<IIQQ", GGUF_MAGIC, GGUF_VERSION, len(self._tensors), len(self._kv))
        with open(self.path, "wb", buffering=16 << 20) as fh:
            fh.write(header)
            fh.write(kv)
            fh.write(infos)
            pos = len(header) + len(kv) + len(infos)
            fh.write(b"\0" * (self._pad(pos) - pos))
            for (name, shape, ttype, producer), nbytes in zip(self._tensors, sizes):
                payload = np.ascontiguousarray(producer()).view(np.uint8).reshape(-1)
                if payload.size != nbytes:
                    raise ValueError(f"{name}: producer returned {payload.size} bytes, expected {nbytes}")
                fh.write(memoryview(payload))
                fh.write(b"\0" * (self._pad(nbytes) - nbytes))
        return self.path
