"""CPU, world_size 2, gloo: the N>1 path — KV-head sharding with no data-path collective, then the
o_proj-shaped all-reduce. Each rank runs the oracle on its head shard; the gathered result must equal
the single-rank result bit for bit per head."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import oracle
from tests.helpers import make_case
from vllm_omni_b200 import tp


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, total_q: int, total_kv: int, q_out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(1)
        c = make_case([33, 100, 7], hq=total_q, hkv=total_kv, seed=11)
        sh = tp.shard_heads(total_q, total_kv, world, rank)
        hs = slice(sh.kv_head_start, sh.kv_head_start + sh.num_kv_heads)
        qs = slice(sh.q_head_start, sh.q_head_start + sh.num_q_heads)
        kc, vc = c.kv_cache[0][:, :, hs], c.kv_cache[1][:, :, hs]
        out, _ = oracle.ref_paged_decode(c.q[:, qs].contiguous(), kc, vc, c.block_table, c.seq_lens, c.scale)
        # attention itself needs no exchange; gather only to compare with the 1-rank run
        parts = [torch.empty_like(out) for _ in range(world)]
        dist.all_gather(parts, out)
        full = torch.cat(parts, dim=1)
        # o_proj stand-in: row-parallel partial products summed by all-reduce
        g = torch.Generator().manual_seed(5)
        w = torch.randn(total_q * 128, 64, generator=g)
        partial = out.reshape(3, -1) @ w[qs.start * 128: qs.stop * 128]
        dist.all_reduce(partial)
        if rank == 0:
            ref, _ = oracle.ref_paged_decode(c.q, c.kv_cache[0], c.kv_cache[1], c.block_table, c.seq_lens, c.scale)
            q_out.put((torch.equal(full, ref), float((partial - ref.reshape(3, -1) @ w).abs().max())))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("total_q,total_kv", [(8, 2), (8, 1)])  # sharded KV heads / replicated KV head
def test_kv_head_sharding_world_size_2(total_q, total_kv):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, total_q, total_kv, q)) for r in range(2)]
    for p in procs:
        p.start()
    same, err = q.get(timeout=180)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert same, "gathered per-rank outputs differ from the single-rank run"
    assert err < 1e-3


def _split_worker(rank: int, world: int, port: int, batch: int, q_out):
    """tp = 2 over ONE KV head (replicated): batch split inside the replica group vs plain replication."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(1)
        total_q, total_kv = 8, 1
        lens = [33, 100, 7, 64, 1][:batch]
        c = make_case(lens, hq=total_q, hkv=total_kv, seed=3)
        sh = tp.shard_heads(total_q, total_kv, world, rank)
        assert sh.kv_replicas == 2 and sh.num_kv_heads == 1
        groups = [dist.new_group(r) for r in tp.ReplicaSplit.groups(world, sh.kv_replicas)]
        rs = tp.ReplicaSplit(sh, groups[rank // sh.kv_replicas])
        per, b0, b1 = rs.plan(batch)
        qs = slice(sh.q_head_start, sh.q_head_start + sh.num_q_heads)
        q_rank = c.q[:, qs].contiguous()
        # reference rule: every rank reads the whole KV head for its own heads
        rep, _ = oracle.ref_paged_decode(q_rank, c.kv_cache[0], c.kv_cache[1], c.block_table, c.seq_lens, c.scale)
        # split: all heads of the group, local sequences only
        q_grp = rs.gather_q(q_rank)
        n = b1 - b0
        assert q_grp.shape == (per, total_q, 128)
        assert torch.equal(q_grp[:n], c.q[b0:b1])
        out_local = torch.zeros(per, total_q, 128)
        if n > 0:
            o, _ = oracle.ref_paged_decode(q_grp[:n], c.kv_cache[0], c.kv_cache[1], c.block_table[b0:b1],
                                           c.seq_lens[b0:b1], c.scale)
            out_local[:n] = o
        mine = rs.scatter_out(out_local, batch)
        q_out.put((rank, bool(torch.equal(mine, rep)), tuple(mine.shape)))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("batch", [4, 5, 1])  # even split, padded split, fewer sequences than ranks
def test_replica_group_batch_split_equals_replication(batch):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_split_worker, args=(r, 2, port, batch, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in range(2)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, same, shape in res:
        assert same, f"rank {rank}: batch-split output differs from the replicated run"
        assert shape == (batch, 4, 128)


def _peer_setup_worker(rank: int, world: int, port: int, q_out):
    """No GPU here: every transport must fail, and every rank must learn that in the same round of the vote."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from vllm_omni_b200 import peer
        try:
            peer.PeerMemory.create(None, 4096, device="cpu")
            res = "created"
        except peer.PeerSetupError as e:
            res = str(e)
        q_out.put((rank, res))
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_peer_memory_setup_failure_is_agreed_by_all_ranks():
    """bench.py falls back to NCCL only if EVERY rank gives up on peer memory: `PeerMemory.create` votes after each
    transport, so a rank whose mapping failed cannot leave the others waiting inside a peer kernel. On this CPU box
    both transports fail; both ranks must raise PeerSetupError naming both ('symm: ... | ipc: ...')."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_peer_setup_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=180) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert set(got) == {0, 1}
    for r, msg in got.items():
        assert msg != "created" and "symm:" in msg and "ipc:" in msg and " | " in msg, (r, msg)
    assert got[0] == got[1]
