"""GGUF container constants.

Format facts (magic, version 3, default alignment 32, 13 KV value types, tensor-type ids and
block geometry) follow the public GGUF specification; parity points in the reference are
`llama.cpp/ggml/include/gguf.h:1-61` (container) and `llama.cpp/ggml/src/ggml-common.h:170-336`
(block structs).  Nothing here is executable reference code - it is the on-disk contract.
"""
from __future__ import annotations

import enum

import numpy as np

GGUF_MAGIC = 0x46554747  # "GGUF" little endian
GGUF_VERSION = 3
GGUF_DEFAULT_ALIGNMENT = 32


class ValueType(enum.IntEnum):
    UINT8 = 0
    INT8 = 1
    UINT16 = 2
    INT16 = 3
    UINT32 = 4
    INT32 = 5
    FLOAT32 = 6
    BOOL = 7
    STRING = 8
    ARRAY = 9
    UINT64 = 10
    INT64 = 11
    FLOAT64 = 12


SCALAR_DTYPES = {
    ValueType.UINT8: np.dtype("<u1"),
    ValueType.INT8: np.dtype("<i1"),
    ValueType.UINT16: np.dtype("<u2"),
    ValueType.INT16: np.dtype("<i2"),
    ValueType.UINT32: np.dtype("<u4"),
    ValueType.INT32: np.dtype("<i4"),
    ValueType.FLOAT32: np.dtype("<f4"),
    ValueType.BOOL: np.dtype("?"),
    ValueType.UINT64: np.dtype("<u8"),
    ValueType.INT64: np.dtype("<i8"),
    ValueType.FLOAT64: np.dtype("<f8"),
}


class GGMLType(enum.IntEnum):
    F32 = 0
    F16 = 1
    Q4_0 = 2
    Q4_1 = 3
    Q5_0 = 6
    Q5_1 = 7
    Q8_0 = 8
    Q8_1 = 9
    Q2_K = 10
    Q3_K = 11
    Q4_K = 12
    Q5_K = 13
    Q6_K = 14
    Q8_K = 15
    IQ2_XXS = 16
    IQ2_XS = 17
    IQ3_XXS = 18
    IQ1_S = 19
    IQ4_NL = 20
    IQ3_S = 21
    IQ2_S = 22
    IQ4_XS = 23
    I8 = 24
    I16 = 25
    I32 = 26
    I64 = 27
    F64 = 28
    IQ1_M = 29
    BF16 = 30
    TQ1_0 = 34
    TQ2_0 = 35
    MXFP4 = 39


QK_K = 256

# type -> (elements per block, bytes per block)
BLOCK_GEOMETRY = {
    GGMLType.F32: (1, 4),
    GGMLType.F16: (1, 2),
    GGMLType.BF16: (1, 2),
    GGMLType.F64: (1, 8),
    GGMLType.I8: (1, 1),
    GGMLType.I16: (1, 2),
    GGMLType.I32: (1, 4),
    GGMLType.I64: (1, 8),
    GGMLType.Q4_0: (32, 2 + 16),
    GGMLType.Q4_1: (32, 4 + 16),
    GGMLType.Q5_0: (32, 2 + 4 + 16),
    GGMLType.Q5_1: (32, 4 + 4 + 16),
    GGMLType.Q8_0: (32, 2 + 32),
    GGMLType.Q8_1: (32, 4 + 32),
    GGMLType.Q2_K: (256, 2 + 2 + 16 + 64),
    GGMLType.Q3_K: (256, 2 + 64 + 32 + 12),
    GGMLType.Q4_K: (256, 2 + 2 + 12 + 128),
    GGMLType.Q5_K: (256, 2 + 2 + 12 + 32 + 128),
    GGMLType.Q6_K: (256, 128 + 64 + 16 + 2),
    GGMLType.Q8_K: (256, 4 + 256 + 32),
    GGMLType.IQ4_NL: (32, 2 + 16),
    GGMLType.IQ4_XS: (256, 2 + 2 + 4 + 128),
    GGMLType.TQ1_0: (256, 48 + 4 + 2),
    GGMLType.TQ2_0: (256, 64 + 2),
    GGMLType.MXFP4: (32, 1 + 16),
    GGMLType.IQ2_XXS: (256, 2 + 64),
    GGMLType.IQ2_XS: (256, 2 + 64 + 8),
    GGMLType.IQ2_S: (256, 2 + 64 + 8 + 8),
    GGMLType.IQ3_XXS: (256, 2 + 96),
    GGMLType.IQ3_S: (256, 2 + 64 + 8 + 32 + 4),
    GGMLType.IQ1_S: (256, 2 + 32 + 16),
    GGMLType.IQ1_M: (256, 32 + 16 + 8),
}


def row_bytes(ggml_type: GGMLType, n_elements: int) -> int:
    """Bytes occupied by `n_elements` contiguous elements of `ggml_type`."""
    per, nbytes = BLOCK_GEOMETRY[GGMLType(ggml_type)]
    if n_elements % per:
        raise ValueError(f"{n_elements} elements is not a multiple of block size {per} for {ggml_type!r}")
    return n_elements // per * nbytes


# File-type ids written to general.file_type (llama_ftype); only those we emit.
FTYPE = {"F32": 0, "F16": 1, "Q4_0": 2, "Q4_1": 3, "Q8_0": 7, "Q5_0": 8, "Q5_1": 9, "Q2_K": 10, "Q3_K": 12, "Q4_K_M": 15,
         "Q5_K_M": 17, "Q6_K": 18, "IQ4_NL": 25, "IQ4_XS": 30, "BF16": 32, "TQ1_0": 36, "TQ2_0": 37, "MXFP4_MOE": 38}


class TokenType(enum.IntEnum):
    NORMAL = 1
    UNKNOWN = 2
    CONTROL = 3
    USER_DEFINED = 4
    UNUSED = 5
    BYTE = 6
