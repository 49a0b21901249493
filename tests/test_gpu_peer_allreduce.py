"""GPU: the peer-memory all-reduce kernels against the sequential fp32 oracle, bit for bit.

The cross-rank flag protocol is exercised on ONE device: `PeerAllReduce.local_ranks(world)` builds `world`
'ranks' whose peer regions are plain allocations on cuda:0, each launched on its own stream, so the
kernels wait for each other exactly as ranks on different GPUs do. (The real multi-process transports are
covered by scripts/peer_check.py under torchrun on 2, 4 and 8 GPUs and by `bench.py --gpus N`.)"""
import pytest
import torch

from oracle import oracle
from vllm_omni_b200 import _lib, peer

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def soft_timeout(monkeypatch):
    """A wait that times out (co-residency not granted on this device) must fail the test, not kill the context."""
    monkeypatch.setenv("B200ATTN_PEER_SOFT_TIMEOUT", "1")


def run_ranks(ars, xs, algo, repeats=1):
    streams = [torch.cuda.Stream() for _ in ars]
    ins = [ar.input(x.shape) for ar, x in zip(ars, xs)]
    for i, x in zip(ins, xs):
        i.copy_(x)
    outs = [torch.full_like(x, float("nan")) for x in xs]
    torch.cuda.synchronize()
    for _ in range(repeats):
        for ar, s, i, o in zip(ars, streams, ins, outs):
            with torch.cuda.stream(s):
                ar.all_reduce(i, out=o, algo=algo)
    torch.cuda.synchronize()
    return outs


@pytest.mark.parametrize("world", [1, 2, 4, 8])
@pytest.mark.parametrize("algo", [_lib.AR_ONESHOT, _lib.AR_TWOSHOT, _lib.AR_PUSH, _lib.AR_PUSH2])
@pytest.mark.parametrize("numel", [8, 8 * 37, 64 * 2048, 128 * 2048 + 24])
def test_allreduce_matches_oracle_bit_exact(lib, world, algo, numel):
    g = torch.Generator().manual_seed(world * 1000 + numel)
    xs_cpu = [(torch.randn(numel, generator=g) * 3).to(torch.bfloat16) for _ in range(world)]
    ars = peer.PeerAllReduce.local_ranks(world, numel, timeout_s=5.0, max_blocks=16)
    if algo in (_lib.AR_PUSH, _lib.AR_PUSH2) and world == 1:
        pytest.skip("PUSH needs peers")
    if numel > 64:                                  # -0.0 is the PUSH sentinel: inputs that contain it must still sum right
        for x in xs_cpu:
            x.view(torch.int16)[5::17] = -32768
    outs = run_ranks(ars, [x.cuda() for x in xs_cpu], algo, repeats=4)   # epochs / the 3 receive buffers carry over launches
    ref = oracle.ref_all_reduce_bf16(xs_cpu)
    for r, o in enumerate(outs):
        assert torch.equal(o.cpu().view(torch.int16), ref.view(torch.int16)), f"rank {r} differs"
        assert ars[r].error_words() == [0, 0, 0, 0]


def test_allreduce_auto_and_graph_replay(lib):
    """AUTO schedule inside CUDA graphs (one per rank), replayed: the flag epochs live in device memory."""
    world, shape = 4, (128, 2048)
    g = torch.Generator().manual_seed(7)
    xs_cpu = [torch.randn(shape, generator=g).to(torch.bfloat16) for _ in range(world)]
    ars = peer.PeerAllReduce.local_ranks(world, shape[0] * shape[1], timeout_s=5.0, max_blocks=16)
    xs = [x.cuda() for x in xs_cpu]
    outs = run_ranks(ars, xs, _lib.AR_AUTO)
    ref = oracle.ref_all_reduce_bf16(xs_cpu)
    assert all(torch.equal(o.cpu(), ref) for o in outs)
    assert ars[0].last_algo == peer.choose_algo(2 * xs[0].numel(), world, False, ars[0].push_slot_bytes) == _lib.AR_PUSH
    graphs, gouts = [], []
    for ar in ars:
        o = torch.empty(shape, dtype=torch.bfloat16, device="cuda")
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            ar.all_reduce(ar.input(shape), out=o)
        graphs.append(gr)
        gouts.append(o)
    torch.cuda.synchronize()
    streams = [torch.cuda.Stream() for _ in ars]
    for it in range(3):
        for ar, x in zip(ars, xs_cpu):
            ar.input(shape).copy_((x.float() * (it + 1)).to(torch.bfloat16))
        torch.cuda.synchronize()
        for gr, s in zip(graphs, streams):
            with torch.cuda.stream(s):
                gr.replay()
        torch.cuda.synchronize()
        ref = oracle.ref_all_reduce_bf16([(x.float() * (it + 1)).to(torch.bfloat16) for x in xs_cpu])
        assert all(torch.equal(o.cpu(), ref) for o in gouts)


def test_allreduce_rejects_bad_arguments(lib):
    ar = peer.PeerAllReduce.local_ranks(1, 64)[0]
    with pytest.raises(ValueError):
        ar.all_reduce(torch.zeros(12, dtype=torch.bfloat16, device="cuda"))      # not a multiple of 8
    with pytest.raises(ValueError):
        ar.all_reduce(torch.zeros(16, dtype=torch.float16, device="cuda"))
    with pytest.raises(ValueError):
        ar.all_reduce(torch.zeros(16, dtype=torch.bfloat16))                     # CPU tensor: no CPU path
    with pytest.raises(_lib.B200AttnError):
        ar.all_reduce(ar.input((16,)), out=ar.input((16,)))                      # out aliases the input region
    with pytest.raises(peer.PeerSetupError):
        ar.all_reduce(ar.input((16,)), algo=_lib.AR_MULTIMEM)


@pytest.mark.parametrize("R,batch", [(2, 64), (2, 7), (4, 10), (8, 33)])
def test_peer_exchange_matches_replica_split_indexing(lib, R, batch):
    """gather_q / scatter_out through peer memory == the all_to_all semantics of tp.ReplicaSplit
    (q_grp[i, g*hq + h] = q_g[b0 + i, h]; out_g[b0 + i, h] = out_grp[i, g*hq + h]), bit for bit, ragged tail
    (batch not divisible by R) included, twice in a row (epochs carry over)."""
    hq, d = 4, 128
    g = torch.Generator().manual_seed(R * 100 + batch)
    qs = [torch.randn(batch, hq, d, generator=g).to(torch.bfloat16) for _ in range(R)]
    exs = peer.PeerExchange.local_ranks(R, batch, hq, d, timeout_s=5.0)
    streams = [torch.cuda.Stream() for _ in exs]
    for ex, q in zip(exs, qs):
        ex.q.copy_(q.cuda())
    torch.cuda.synchronize()
    # warm-up, phase by phase: the FIRST launch of a kernel loads its module, which can block the host behind a
    # running kernel — with all 'ranks' driven by one host thread that would stall rank 1's launch behind rank 0's
    # spinning kernel (a single-process artefact; real ranks have a host thread each)
    for ex in exs:                                   # torch's own kernels first, with nothing spinning
        ex.out_grp.copy_(ex.q_grp)
        ex.out_grp.mul_(1)
    torch.cuda.synchronize()
    for phase in ("gather_q", "scatter_out"):
        for ex, s in zip(exs, streams):
            with torch.cuda.stream(s):
                getattr(ex, phase)()
        torch.cuda.synchronize()
    for it in range(2):
        for ex, s in zip(exs, streams):
            with torch.cuda.stream(s):
                ex.gather_q()
                ex.out_grp.copy_(ex.q_grp)          # stand-in for the decode; in place: a torch allocation here could
                ex.out_grp.mul_(2 + it)              # block the host behind rank 0's kernel, which waits for rank 1's launch
                ex.scatter_out()
        torch.cuda.synchronize()
        for i, ex in enumerate(exs):
            n = ex.b1 - ex.b0
            want = torch.stack([q[ex.b0:ex.b1] for q in qs], dim=1).reshape(n, R * hq, d)
            assert torch.equal(ex.q_grp[:n].cpu(), want), f"gather differs on rank {i}"
            assert torch.equal(ex.out.cpu(), qs[i].float().mul(2 + it).to(torch.bfloat16)), f"scatter differs on rank {i}"
            off = peer.ERROR_WORDS_OFFSET
            assert ex.mem.local[off: off + 16].view(torch.int32).tolist() == [0, 0, 0, 0]


def test_exchange_serves_smaller_live_batches(lib):
    """A serving process sizes the exchange once (max_num_seqs) and then decodes batches of any smaller size: the
    rows, the split and the padding follow the LIVE batch (all ranks pass the same one)."""
    R, hq, d, cap = 2, 4, 128, 11
    exs = peer.PeerExchange.local_ranks(R, cap, hq, d, timeout_s=5.0)
    streams = [torch.cuda.Stream() for _ in exs]
    for ex in exs:                                   # load torch's kernels with nothing spinning (see above)
        ex.out_grp.copy_(ex.q_grp)
        ex.out_grp.mul_(1)
    torch.cuda.synchronize()
    with pytest.raises(ValueError):
        exs[0].split(cap + 1)
    for it, n in enumerate((cap, 5, 8, 1)):
        g = torch.Generator().manual_seed(1000 + n)
        qs = [torch.randn(n, hq, d, generator=g).to(torch.bfloat16) for _ in range(R)]
        for ex, q in zip(exs, qs):
            ex.q[:n].copy_(q.cuda())
        torch.cuda.synchronize()
        for phase in (0, 1):
            for ex, s in zip(exs, streams):
                with torch.cuda.stream(s):
                    if phase == 0:
                        ex.gather_q(n)
                    else:
                        ex.out_grp.copy_(ex.q_grp)
                        ex.out_grp.mul_(3)
                        ex.scatter_out(n)
            torch.cuda.synchronize()
        for i, ex in enumerate(exs):
            per, b0, b1 = ex.split(n)
            want = torch.stack([q[b0:b1] for q in qs], dim=1).reshape(b1 - b0, R * hq, d)
            assert torch.equal(ex.q_grp[:b1 - b0].cpu(), want), f"gather differs on rank {i} at batch {n}"
            assert torch.equal(ex.out[:n].cpu(), qs[i].float().mul(3).to(torch.bfloat16)), f"scatter, rank {i}, batch {n}"
