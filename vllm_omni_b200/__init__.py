"""B200-native paged-attention decode path for vllm-omni's autoregressive stages.

Package contents (only what the hot path needs):
  csrc/        hand-written sm_100a CUDA kernels + the C ABI (include/b200attn.h)
  _lib.py      ctypes binding of libb200attn.so (fails loudly if the library is missing)
  ops.py       tensor-level wrappers over the C ABI
  layout.py    KV-cache layout contract
  tp.py        KV-head tensor-parallel sharding rule of the reference
  backend.py   vLLM AttentionBackend / MetadataBuilder / AttentionImpl (imports vllm)
  plugin.py    registration as AttentionBackendEnum.CUSTOM
  workloads.py named decode workloads + synthetic inputs
"""
__version__ = "0.1.0"
