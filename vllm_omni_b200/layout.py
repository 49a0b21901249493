"""KV-cache layout contract of the B200 backend (no vLLM import, usable from tests and bench).

Logical shape  (what vLLM and the reference's KV-transfer reader index):
    kv_cache[2, num_blocks, block_size, num_kv_heads, head_size]
    — `kv_cache[0][block_id]` is a [block_size, Hkv, d] key block, as required by
      vllm_omni/distributed/omni_connectors/utils/kv_utils.py:417-450 and
      kv_transfer_manager.py:829-846.
Physical order (what the kernels stream):
    [num_blocks, num_kv_heads, 2, block_size, head_size]
    — for one (block, kv head) the K page and the V page are adjacent: block_size*d*2 elements
      (8 KiB at bs=16, d=128, bf16) of contiguous HBM per work unit, each page one TMA box row
      range. Declared to vLLM through `get_kv_cache_stride_order` = (1, 3, 0, 2, 4).
"""
from __future__ import annotations

import torch

KV_CACHE_STRIDE_ORDER = (1, 3, 0, 2, 4)


def kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int, head_size: int) -> tuple[int, ...]:
    if block_size % 16 != 0:
        raise ValueError("block_size must be a multiple of 16")
    return (2, num_blocks, block_size, num_kv_heads, head_size)


def cache_torch_dtype(kv_cache_dtype: str) -> torch.dtype:
    return torch.uint8 if kv_cache_dtype in ("fp8", "fp8_e4m3") else torch.bfloat16


def alloc_kv_cache(num_blocks: int, block_size: int, num_kv_heads: int, head_size: int,
                   kv_cache_dtype: str = "auto", device: torch.device | str = "cuda",
                   stride_order: tuple[int, ...] = KV_CACHE_STRIDE_ORDER,
                   zero: bool = True) -> torch.Tensor:
    """Allocate a cache the way vLLM's runner does for this backend: a flat buffer viewed in the
    physical order and permuted back to the logical shape (vllm/v1/worker/gpu_model_runner.py:6935-6960)."""
    shape = kv_cache_shape(num_blocks, block_size, num_kv_heads, head_size)
    phys = tuple(shape[i] for i in stride_order)
    inv = [stride_order.index(i) for i in range(len(stride_order))]
    dt = cache_torch_dtype(kv_cache_dtype)
    raw = (torch.zeros if zero else torch.empty)(phys, dtype=dt, device=device)
    return raw.permute(*inv)
