#!/usr/bin/env python3
"""Regenerate `distributed_llm_pipeline_b200/gguf/iq_grids.npz`: the code books of the IQ1 / IQ2 / IQ3 tensor types.

The code books are constants of the file format (reference: the `iq2xxs_grid` ... `iq1s_grid` tables of
`llama.cpp/ggml/src/ggml-common.h`); the installed `gguf` package carries the same tables, and this script reads them from there.
Each book is stored as small integers (the lattice coordinates themselves), one row per code.
"""
import os
import sys

import numpy as np


def main() -> int:
    import gguf.quants as q
    out = {}
    for name in ("IQ2_XXS", "IQ2_XS", "IQ2_S", "IQ3_XXS", "IQ3_S", "IQ1_S"):
        cls = getattr(q, name)
        cls.init_grid()
        g = np.asarray(cls.grid, np.float32).reshape(cls.grid_shape)
        assert np.array_equal(g, np.round(g)) and np.abs(g).max() < 128
        out[name.lower()] = g.astype(np.int8)
    dst = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "distributed_llm_pipeline_b200", "gguf", "iq_grids.npz")
    np.savez_compressed(dst, **out)
    print({k: v.shape for k, v in out.items()}, os.path.getsize(dst), "bytes")
    return 0


if __name__ == "__main__":
    sys.exit(main())
