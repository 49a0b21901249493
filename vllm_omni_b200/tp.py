"""KV-head tensor-parallel sharding rule of the AR stages.

Mirrors the reference's Qwen2Attention.__init__
(vllm_omni/model_executor/models/qwen2_5_omni/qwen2_old.py:89-106; vLLM qwen3_moe.py:279-292):
query heads are split evenly; KV heads are split when total_kv >= tp and replicated when
tp > total_kv (each KV head then lives on tp / total_kv ranks). Per-head FP8 scales follow the
KV heads (vllm compressed_tensors.py:977-988).
"""
from __future__ import annotations

from dataclasses import dataclass


@dataclass(frozen=True)
class HeadShard:
    tp_size: int
    rank: int
    num_q_heads: int          # per rank
    num_kv_heads: int         # per rank (>= 1)
    q_head_start: int         # first global query head of this rank
    kv_head_start: int        # first global KV head of this rank
    kv_replicas: int          # ranks holding the same KV head (1 = sharded)

    @property
    def q_heads(self) -> range:
        return range(self.q_head_start, self.q_head_start + self.num_q_heads)

    @property
    def kv_heads(self) -> range:
        return range(self.kv_head_start, self.kv_head_start + self.num_kv_heads)


def shard_heads(total_q_heads: int, total_kv_heads: int, tp_size: int, rank: int) -> HeadShard:
    if tp_size < 1 or not 0 <= rank < tp_size:
        raise ValueError(f"bad tp_size/rank {tp_size}/{rank}")
    if total_q_heads % tp_size != 0:
        raise ValueError(f"{total_q_heads} query heads not divisible by tp {tp_size}")
    hq = total_q_heads // tp_size
    if total_kv_heads >= tp_size:
        if total_kv_heads % tp_size != 0:          # qwen2_old.py:95-97
            raise ValueError(f"{total_kv_heads} KV heads not divisible by tp {tp_size}")
        hkv = total_kv_heads // tp_size
        return HeadShard(tp_size, rank, hq, hkv, rank * hq, rank * hkv, 1)
    if tp_size % total_kv_heads != 0:              # qwen2_old.py:98-101
        raise ValueError(f"tp {tp_size} not a multiple of {total_kv_heads} KV heads")
    rep = tp_size // total_kv_heads
    return HeadShard(tp_size, rank, hq, 1, rank * hq, rank // rep, rep)


def shard_scales(scales, shard: HeadShard):
    """Slice a per-KV-head scale tensor [total_kv] for this rank (per-tensor [1] passes through)."""
    if scales is None or scales.numel() == 1:
        return scales
    return scales[shard.kv_head_start: shard.kv_head_start + shard.num_kv_heads].contiguous()
