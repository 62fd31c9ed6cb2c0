"""GGUF container I/O, block-quant codecs and synthetic checkpoint writer."""
from .constants import BLOCK_GEOMETRY, GGMLType, ValueType, row_bytes
from .quants import dequantize, quantize, random_blocks
from .reader import GGUFFile, GGUFModelFiles, TensorInfo, open_gguf
from .synth import synthetic_vocab, tensor_type_for, write_synthetic_gguf
from .writer import GGUFWriter

__all__ = ["BLOCK_GEOMETRY", "GGMLType", "ValueType", "row_bytes", "dequantize", "quantize", "random_blocks",
           "GGUFFile", "GGUFModelFiles", "TensorInfo", "open_gguf", "synthetic_vocab", "tensor_type_for",
           "write_synthetic_gguf", "GGUFWriter"]
