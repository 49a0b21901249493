"""Named decode workloads (BASELINE.json `configs`) and their synthetic inputs.

Shapes follow SURVEY.md §8: per-rank head counts come from the reference's TP rule (`tp.py`),
pages are NON-contiguous (block ids are a seeded random permutation), the KV history is written
into the paged cache through the cache-write kernel itself, and the new token of every sequence
sits at position S_b - 1. Seeds are fixed so every rank and every run sees the same tensors.
"""
from __future__ import annotations

from dataclasses import dataclass, field, replace

import torch

from . import layout
from .tp import HeadShard, shard_heads


@dataclass(frozen=True)
class Workload:
    name: str
    model: str
    total_q_heads: int
    total_kv_heads: int
    head_size: int
    num_layers: int
    hidden_size: int
    batch: int
    ctx: int
    kv_cache_dtype: str = "auto"   # "auto" (bf16) or "fp8_e4m3"
    block_size: int = 16

    @property
    def fp8(self) -> bool:
        return self.kv_cache_dtype in ("fp8", "fp8_e4m3")

    @property
    def kv_elem_bytes(self) -> int:
        return 1 if self.fp8 else 2

    def kv_bytes_per_layer(self, num_kv_heads: int | None = None, total_tokens: int | None = None) -> int:
        """Algorithmic bytes of one decode step of one layer: K and V of every context token once
        (sum_b S_b * 2 * Hkv * d * sizeof(kv))."""
        hkv = self.total_kv_heads if num_kv_heads is None else num_kv_heads
        toks = self.batch * self.ctx if total_tokens is None else total_tokens
        return toks * 2 * hkv * self.head_size * self.kv_elem_bytes


# Layer counts / hidden sizes: Qwen2.5-Omni-7B from the transformers class defaults; the others
# from the public HF config.json of the checkpoints (not in the reference tree; SURVEY.md §8 †).
WORKLOADS: dict[str, Workload] = {
    "cfg1": Workload("cfg1", "Qwen2.5-Omni-3B thinker", 16, 2, 128, 36, 2048, 1, 512),
    "cfg2": Workload("cfg2", "Qwen2.5-Omni-7B thinker", 28, 4, 128, 28, 3584, 32, 2048),
    "cfg3": Workload("cfg3", "Qwen3-Omni-30B-A3B thinker", 32, 4, 128, 48, 2048, 64, 4096),
    "cfg4": Workload("cfg4", "Qwen3-Omni-30B-A3B thinker (fp8 KV)", 32, 4, 128, 48, 2048, 128, 8192,
                     "fp8_e4m3"),
    "cfg5": Workload("cfg5", "Qwen3-Omni-30B-A3B thinker (32k)", 32, 4, 128, 48, 2048, 16, 32768),
}

SEED_Q, SEED_KV, SEED_BLOCKS, SEED_LENS, SEED_NEW = 0, 1, 2, 3, 4


def make_seq_lens(batch: int, ctx: int, ragged: bool = False) -> torch.Tensor:
    """Uniform `ctx`, or U[ctx/2, ctx] (seed 3) for the ragged variant. CPU int32."""
    if not ragged:
        return torch.full((batch,), ctx, dtype=torch.int32)
    g = torch.Generator().manual_seed(SEED_LENS)
    lo = max(1, ctx // 2)
    return torch.randint(lo, ctx + 1, (batch,), generator=g, dtype=torch.int32)


def make_block_table(seq_lens: torch.Tensor, block_size: int, spare_blocks: int = 64,
                     contiguous: bool = False) -> tuple[torch.Tensor, int]:
    """Block table [B, max_blocks] (CPU int32) over a pool of sum(pages)+spare blocks; block ids are
    a seeded random permutation of the pool (seed 2) unless `contiguous`. Unused entries are 0."""
    pages = (seq_lens.to(torch.int64) + block_size - 1) // block_size
    total = int(pages.sum())
    nb = total + spare_blocks
    if contiguous:
        perm = torch.arange(nb, dtype=torch.int32)
    else:
        g = torch.Generator().manual_seed(SEED_BLOCKS)
        perm = torch.randperm(nb, generator=g).to(torch.int32)
    max_blocks = max(1, int(pages.max()) if pages.numel() else 1)
    bt = torch.zeros((seq_lens.numel(), max_blocks), dtype=torch.int32)
    off = 0
    for b, n in enumerate(pages.tolist()):
        bt[b, :n] = perm[off:off + n]
        off += n
    return bt, nb


def slots_of_positions(block_table: torch.Tensor, seq_idx: torch.Tensor, pos: torch.Tensor,
                       block_size: int) -> torch.Tensor:
    """slot = block_table[b, pos // bs] * bs + pos % bs  (int64)."""
    blk = block_table[seq_idx.long(), (pos // block_size).long()].to(torch.int64)
    return blk * block_size + (pos % block_size).to(torch.int64)


def default_scales(num_kv_heads_total: int, device="cpu") -> tuple[torch.Tensor, torch.Tensor]:
    """Static per-KV-head FP8 scales for N(0,1) data: absmax ~ 6 mapped near the e4m3 maximum,
    made distinct per head so a per-head indexing slip cannot hide."""
    h = torch.arange(num_kv_heads_total, dtype=torch.float32)
    k = (6.0 / 448.0) * (1.0 + 0.05 * h)
    v = (6.0 / 448.0) * (1.0 + 0.03 * h)
    return k.to(device), v.to(device)


@dataclass
class DecodeInputs:
    workload: Workload
    shard: HeadShard
    q: torch.Tensor                      # [B, Hq_local, d] bf16
    new_k: torch.Tensor                  # [B, Hkv_local, d] bf16  (token at position S_b-1)
    new_v: torch.Tensor
    slot_mapping: torch.Tensor           # [B] int64
    block_table: torch.Tensor            # [B, max_blocks] int32
    seq_lens: torch.Tensor               # [B] int32
    max_seq_len: int
    num_blocks: int
    kv_caches: list[torch.Tensor] = field(default_factory=list)   # per layer, logical [2,nb,bs,Hkv,d]
    k_scale: torch.Tensor | None = None  # [Hkv_local] f32 (fp8 only)
    v_scale: torch.Tensor | None = None
    softmax_scale: float = 0.0

    @property
    def total_tokens(self) -> int:
        return int(self.seq_lens.sum())

    @property
    def kv_bytes_per_layer(self) -> int:
        return self.workload.kv_bytes_per_layer(self.shard.num_kv_heads, self.total_tokens)


def _randn_bf16(shape, seed: int, device) -> torch.Tensor:
    g = torch.Generator(device=device).manual_seed(seed)
    return torch.randn(shape, generator=g, device=device, dtype=torch.float32).to(torch.bfloat16)


def fill_history(kv_cache: torch.Tensor, block_table_dev: torch.Tensor, seq_lens_cpu: torch.Tensor,
                 wl: Workload, shard: HeadShard, layer: int, k_scale, v_scale,
                 chunk_tokens: int = 1 << 18) -> None:
    """Write N(0,1) K/V for positions [0, S_b - 1) of every sequence through the cache-write kernel."""
    from . import ops  # CUDA path only
    dev = kv_cache.device
    k_cache, v_cache = ops.split_kv_cache(kv_cache)
    lens = (seq_lens_cpu.to(torch.int64) - 1).clamp_min(0)
    seq_idx = torch.repeat_interleave(torch.arange(lens.numel()), lens)
    starts = torch.cumsum(lens, 0) - lens
    pos = torch.arange(int(lens.sum())) - torch.repeat_interleave(starts, lens)
    seq_idx, pos = seq_idx.to(dev), pos.to(dev)
    n = pos.numel()
    for c0 in range(0, n, chunk_tokens):
        c1 = min(n, c0 + chunk_tokens)
        slots = slots_of_positions(block_table_dev, seq_idx[c0:c1], pos[c0:c1], wl.block_size)
        # one generator stream per (layer, chunk); heads are sliced from the global-head tensor so
        # every TP rank sees the same values for the heads it owns
        seed = SEED_KV + 7919 * layer + 104729 * (c0 // chunk_tokens)
        kv = _randn_bf16((2, c1 - c0, wl.total_kv_heads, wl.head_size), seed, dev)
        hs = slice(shard.kv_head_start, shard.kv_head_start + shard.num_kv_heads)
        ops.reshape_and_cache(kv[0][:, hs], kv[1][:, hs], k_cache, v_cache, slots,
                              wl.kv_cache_dtype, k_scale, v_scale)


def make_decode_inputs(wl: Workload, device="cuda", *, tp_size: int = 1, rank: int = 0,
                       ragged: bool = False, num_layers: int = 1, contiguous_blocks: bool = False,
                       fill: bool = True) -> DecodeInputs:
    """Synthetic decode step for `wl` on one TP rank, `num_layers` independent KV caches."""
    shard = shard_heads(wl.total_q_heads, wl.total_kv_heads, tp_size, rank)
    seq_lens = make_seq_lens(wl.batch, wl.ctx, ragged)
    bt, nb = make_block_table(seq_lens, wl.block_size, contiguous=contiguous_blocks)
    dev = torch.device(device)
    q_all = _randn_bf16((wl.batch, wl.total_q_heads, wl.head_size), SEED_Q, dev)
    new_all = _randn_bf16((2, wl.batch, wl.total_kv_heads, wl.head_size), SEED_NEW, dev)
    hs = slice(shard.kv_head_start, shard.kv_head_start + shard.num_kv_heads)
    qs = slice(shard.q_head_start, shard.q_head_start + shard.num_q_heads)
    k_scale = v_scale = None
    if wl.fp8:
        ks, vs = default_scales(wl.total_kv_heads, dev)
        k_scale, v_scale = ks[hs].contiguous(), vs[hs].contiguous()
    bt_dev, lens_dev = bt.to(dev), seq_lens.to(dev)
    last = (seq_lens.to(torch.int64) - 1).clamp_min(0)
    slot_mapping = slots_of_positions(bt, torch.arange(wl.batch), last, wl.block_size)
    slot_mapping = torch.where(seq_lens > 0, slot_mapping, torch.full_like(slot_mapping, -1)).to(dev)
    inp = DecodeInputs(
        workload=wl, shard=shard, q=q_all[:, qs].contiguous(), new_k=new_all[0][:, hs].contiguous(),
        new_v=new_all[1][:, hs].contiguous(), slot_mapping=slot_mapping, block_table=bt_dev,
        seq_lens=lens_dev, max_seq_len=int(seq_lens.max()) if wl.batch else 0, num_blocks=nb,
        k_scale=k_scale, v_scale=v_scale, softmax_scale=wl.head_size ** -0.5)
    for layer in range(num_layers):
        cache = layout.alloc_kv_cache(nb, wl.block_size, shard.num_kv_heads, wl.head_size,
                                      wl.kv_cache_dtype, dev)
        if fill:
            fill_history(cache, bt_dev, seq_lens, wl, shard, layer, k_scale, v_scale)
        inp.kv_caches.append(cache)
    return inp


def scaled(wl: Workload, **kw) -> Workload:
    """A copy of `wl` with fields overridden (tests run the named shapes at reduced batch/ctx)."""
    return replace(wl, **kw)
