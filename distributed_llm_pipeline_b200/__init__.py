"""distributed_llm_pipeline_b200 - a Blackwell-native distributed LLM inference engine.

Same capabilities as un1c4on/Distributed-LLM-Pipeline (GGUF loading, layer-split pipeline stages,
row-split tensor parallelism, KV-cached decode, SSE orchestrator + static UI), designed B200-first:
one process per GPU, hand-written sm_100a kernels behind a C ABI (`include/dlp_b200.h`), NCCL for
plumbing and in-kernel NVLink peer traffic for the fused GEMM+collective paths.
"""
__version__ = "0.1.0"
