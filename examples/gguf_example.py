#!/usr/bin/env python
"""gguf: write a small GGUF file with a few key-value pairs and tensors, read it back and check the data.   (llama.cpp/examples/gguf)"""
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from distributed_llm_pipeline_b200.gguf import GGMLType, GGUFWriter, open_gguf, quantize  # noqa: E402
from distributed_llm_pipeline_b200.gguf.constants import ValueType  # noqa: E402

path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(tempfile.mkdtemp(), "example.gguf")
w = GGUFWriter(path)
w.add("some.parameter.uint8", 0x12, ValueType.UINT8)
w.add("some.parameter.int32", -0x12345678, ValueType.INT32)
w.add("some.parameter.float32", 0.123456789, ValueType.FLOAT32)
w.add("some.parameter.bool", True)
w.add("some.parameter.string", "hello world")
w.add("some.parameter.arr.i32", [1, 2, 3, 4], etype=ValueType.INT32)
w.add("some.parameter.arr.str", ["hello", "world", "!"], etype=ValueType.STRING)
rng = np.random.default_rng(0)
tensors = {f"tensor_{i}": (rng.standard_normal((4 + i, 64)) * 0.1).astype(np.float32) for i in range(4)}
for i, (name, arr) in enumerate(tensors.items()):
    t = [GGMLType.F32, GGMLType.F16, GGMLType.Q8_0, GGMLType.Q4_0][i]
    w.add_tensor(name, arr.shape, t, lambda arr=arr, t=t: quantize(arr, t))
w.write()
f = open_gguf(path)
print(f"{path}: {len(f.metadata)} key-value pairs, {len(f.tensors)} tensors")
for k, v in f.metadata.items():
    print(f"  {k} = {v!r}")
ok = True
for name, info in f.tensors.items():
    got = f.array(name)
    err = float(np.abs(got - tensors[name]).max())
    ok &= err < 0.05          # Q4_0 steps are amax / 8 wide
    print(f"  {name}: {info.shape} {GGMLType(info.ggml_type).name}, max |error| vs the written values {err:.2e}")
print("data check passed" if ok else "DATA MISMATCH")
raise SystemExit(0 if ok else 1)
