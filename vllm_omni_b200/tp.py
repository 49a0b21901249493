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


class ReplicaSplit:
    """Batch split inside a KV-head replica group (tp > total_kv_heads).

    The reference rule replicates a KV head on R = tp / total_kv_heads ranks; each of them then
    streams the SAME KV pages of every sequence for its own 1/R of the head group, so the step's
    dominant cost (KV bytes read) is paid R times. The R ranks hold identical caches, so instead each
    rank can serve ALL query heads of the group for 1/R of the sequences: per-rank KV bytes drop by R
    and the group exchanges only queries and outputs (B x Hq_rank x d bf16 each way — 0.4 % of the KV
    bytes at cfg3). Results are identical head for head; the o_proj input layout is unchanged.

    Plumbing is two `all_to_all_single` calls on the replica group (NCCL over NVLink on GPUs, gloo in
    the CPU tests): `gather_q` before the decode, `scatter_out` after it. With `exchange` (a
    `peer.PeerExchange` over the same group) both become one kernel each of this repo over peer memory
    (`b200attn_peer_exchange`): same rows, no NCCL, no staging permutes.
    """

    def __init__(self, shard: HeadShard, group=None, exchange=None):
        import torch.distributed as dist  # local: tp.py stays importable without torch.distributed use

        self.dist = dist
        self.R = shard.kv_replicas
        self.index = shard.rank % self.R          # position inside the replica group
        self.group = group
        self.hq = shard.num_q_heads
        self.exchange = exchange

    @staticmethod
    def groups(tp_size: int, kv_replicas: int):
        """Rank lists of all replica groups (every rank must create every group, in this order)."""
        return [list(range(g * kv_replicas, (g + 1) * kv_replicas)) for g in range(tp_size // kv_replicas)]

    @property
    def group_q_heads(self) -> int:
        return self.R * self.hq

    def plan(self, batch: int) -> tuple[int, int, int]:
        """(sequences per rank incl. padding, first local sequence, one past the last local sequence)."""
        per = -(-batch // self.R)
        b0 = min(batch, self.index * per)
        return per, b0, min(batch, b0 + per)

    def _pad(self, x, rows: int):
        import torch
        if x.shape[0] == rows:
            return x
        pad = torch.zeros((rows - x.shape[0], *x.shape[1:]), dtype=x.dtype, device=x.device)
        return torch.cat([x, pad])

    def gather_q(self, q, out=None):
        """q [B, Hq_rank, d] of this rank -> [per, R * Hq_rank, d]: all heads of the group for the
        local sequences [b0, b1) (rows past b1 - b0 are padding)."""
        import torch
        per, _, _ = self.plan(q.shape[0])
        ex = self.exchange
        if ex is not None:
            n = q.shape[0]                      # the live batch: anything up to the size the exchange was made for
            if q.data_ptr() != ex.q.data_ptr():
                ex.q[:n].copy_(q)
            res = ex.gather_q(n)[:per]
            if out is not None and out.data_ptr() != res.data_ptr():
                out.copy_(res)
                return out
            return res
        send = self._pad(q, self.R * per).contiguous()
        recv = torch.empty_like(send)             # [R, per, hq, d]: chunk i = peer i's heads, my sequences
        self.dist.all_to_all_single(recv, send, group=self.group)
        d = q.shape[-1]
        res = recv.view(self.R, per, self.hq, d).permute(1, 0, 2, 3).reshape(per, self.R * self.hq, d)
        if out is not None:
            out.copy_(res)
            return out
        return res.contiguous()

    def scatter_out(self, out_local, batch: int, out=None):
        """out_local [per, R * Hq_rank, d] -> [B, Hq_rank, d]: this rank's heads for every sequence."""
        import torch
        per, d = out_local.shape[0], out_local.shape[-1]
        ex = self.exchange
        if ex is not None:
            if out_local.data_ptr() != ex.out_grp.data_ptr():
                ex.out_grp[:out_local.shape[0]].copy_(out_local)
            res = ex.scatter_out(batch)[:batch]
            if out is not None and out.data_ptr() != res.data_ptr():
                out.copy_(res)
                return out
            return res
        send = out_local.view(per, self.R, self.hq, d).permute(1, 0, 2, 3).contiguous()
        recv = torch.empty_like(send)             # chunk i = peer i's sequences, my heads
        self.dist.all_to_all_single(recv, send, group=self.group)
        res = recv.view(self.R * per, self.hq, d)[:batch]
        if out is not None:
            out.copy_(res)
            return out
        return res



def replica_split_for_vllm(vllm_config, max_num_seqs: int | None = None):
    """Build the `ReplicaSplit` (+ peer-memory exchange) of THIS worker from vLLM's tensor-parallel group — the place
    the reference initialises TP is `GPUARWorker.init_device` (vllm_omni/worker/gpu_ar_worker.py:69-75, vLLM's
    `init_worker_distributed_environment`); this is called afterwards, from the metadata builder, once per worker.
    Returns None when no KV head is replicated (tp <= total_kv_heads) or the opt-in `B200ATTN_REPLICA_SPLIT` is off:
    the reference rule (every rank of a replica group streams the whole KV head) stays the default."""
    import os
    if os.environ.get("B200ATTN_REPLICA_SPLIT", "0") in ("", "0"):
        return None
    from vllm.distributed import get_tensor_model_parallel_rank, get_tp_group
    mc, pc = vllm_config.model_config, vllm_config.parallel_config
    tp = pc.tensor_parallel_size
    total_kv = mc.get_total_num_kv_heads()
    if tp <= total_kv:
        return None
    import torch.distributed as dist
    rank = get_tensor_model_parallel_rank()
    shard = shard_heads(mc.hf_text_config.num_attention_heads, total_kv, tp, rank)
    tp_ranks = get_tp_group().ranks                       # global ranks of this TP group, in TP order
    my_group = None
    for grp in ReplicaSplit.groups(tp, shard.kv_replicas):  # every rank creates every group, in the same order
        g = dist.new_group([tp_ranks[i] for i in grp])
        if rank in grp:
            my_group = g
    exchange = None
    try:
        from . import peer
        n = max_num_seqs or vllm_config.scheduler_config.max_num_seqs
        exchange = peer.PeerExchange.create(my_group, n, shard.num_q_heads, mc.get_head_size())
    except Exception:  # noqa: BLE001 - NCCL all_to_all is the documented fallback of ReplicaSplit
        exchange = None
    return ReplicaSplit(shard, my_group, exchange=exchange)
