"""Peer-memory all-reduce for the o_proj row-parallel sum that follows the attention path.

Host side of `b200attn_peer_allreduce` (include/b200attn.h; kernels in csrc/peer_allreduce.cu): one
process per GPU, every rank owns one peer-mapped allocation `[flags | input | scratch/out]`, and the
all-reduce is ONE kernel launch that reads the peers' input regions over NVLink (or lets the NVSwitch
reduce through a multicast mapping). It stands where vLLM calls `tensor_model_parallel_all_reduce`
(vllm/distributed/communication_op.py:12) after `self.o_proj(attn_output)`
(vllm_omni/model_executor/models/qwen2_5_omni/qwen2_old.py:155).

PyTorch is plumbing here: `torch.distributed._symmetric_memory` (or CUDA IPC handles shipped with
`all_gather_object`) provides the peer mappings; the collective itself is this repo's kernel. There
is no fallback inside this module: a transport that cannot be set up raises `PeerSetupError` and the
caller decides (bench.py then reports NCCL and says so).
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib

ALGO_NAMES = {_lib.AR_AUTO: "auto", _lib.AR_ONESHOT: "oneshot", _lib.AR_TWOSHOT: "twoshot",
              _lib.AR_MULTIMEM: "multimem", _lib.AR_PUSH: "push", _lib.AR_PUSH2: "push2"}
PUSH_MAX_BYTES = 1 << 20      # largest message the PUSH schedule is provisioned for (per slot)
_ALIGN = 256
# byte offset of the 4 error words inside a flag region: after [128 blocks][8 ranks] arrival slots and 128 epochs
# (kPeerMaxBlocks / kPeerMaxRanks in csrc/kernels.h)
ERROR_WORDS_OFFSET = (128 * 8 + 128) * 4


class PeerSetupError(RuntimeError):
    """Peer-mapped memory could not be established between the ranks."""


def region_layout(max_numel: int, world: int = 0) -> tuple[int, int, int, int]:
    """(input offset, scratch/out offset, region capacity in bytes, total bytes) of one rank's allocation
    for bf16 tensors of up to `max_numel` elements: [flags | input | scratch/out | push receive buffers]. The
    scratch region doubles as the multimem output region, so it has the full capacity. With `world` > 1 the
    allocation ends with the 3 x world receive slots of the PUSH schedule (see `push_layout`)."""
    cap = -(-(max_numel * 2 + 16) // _ALIGN) * _ALIGN
    off_in = _lib.PEER_FLAG_BYTES
    off_scr = off_in + cap
    return off_in, off_scr, cap, off_scr + cap + 3 * world * push_layout(max_numel, world)


def push_layout(max_numel: int, world: int) -> int:
    """Bytes of one receive slot of the PUSH schedule (0 = not provisioned)."""
    if world <= 1:
        return 0
    return min(-(-(max_numel * 2) // _ALIGN) * _ALIGN, PUSH_MAX_BYTES)


def choose_algo(nbytes: int, world: int, has_multicast: bool, push_slot_bytes: int = 0) -> int:
    """The schedule `AUTO` resolves to (host arithmetic in the C library, no GPU needed)."""
    return int(_lib.load().b200attn_peer_allreduce_algo(nbytes, world, int(bool(has_multicast)), push_slot_bytes))


class _RawCuda:
    """`__cuda_array_interface__` view of raw device memory (lets torch wrap a cudaIpc pointer)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False),
                                         "version": 3, "strides": None}


def _ptr_array(ptrs):
    return (C.c_void_p * len(ptrs))(*[C.c_void_p(int(p)) for p in ptrs])


class PeerMemory:
    """One peer-mapped allocation per rank of `group`, zero-filled: `local` is this rank's bytes as a
    uint8 tensor, `base_ptrs[r]` is rank r's allocation as mapped into this process, `multicast_ptr` the
    multicast mapping of all of them (0 if the fabric has none)."""

    def __init__(self, world, rank, base_ptrs, local, multicast_ptr=0, transport="local", keepalive=None):
        self.world, self.rank = world, rank
        self.base_ptrs = [int(p) for p in base_ptrs]
        self.local = local
        self.multicast_ptr = int(multicast_ptr or 0)
        self.transport = transport
        self._keep = keepalive

    @classmethod
    def create(cls, group=None, nbytes: int = 0, device=None, transport: str = "auto"):
        """Collective over `group` (default: WORLD). transport: 'symm' (torch symmetric memory; gives a
        multicast mapping where the fabric has one), 'ipc' (cudaMalloc + cudaIpc handles) or 'auto'
        (symm, then ipc). Raises PeerSetupError on every rank if any rank could not map its peers."""
        import torch.distributed as dist
        group = group if group is not None else dist.group.WORLD
        device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        errors = []
        for tr in (("symm", "ipc") if transport == "auto" else (transport,)):
            try:
                mem = (cls._create_symm if tr == "symm" else cls._create_ipc)(group, nbytes, device)
            except Exception as ex:  # noqa: BLE001 - every rank must take the same branch: vote below
                mem, err = None, f"{tr}: {type(ex).__name__}: {ex}"
            else:
                err = None
            votes = [None] * dist.get_world_size(group)
            dist.all_gather_object(votes, err, group=group)
            if all(v is None for v in votes):
                torch.cuda.synchronize(device)
                dist.barrier(group)          # every rank's allocation is zeroed before anyone launches
                return mem
            errors.append("; ".join(sorted({v for v in votes if v})))
        raise PeerSetupError(" | ".join(errors))

    @classmethod
    def _create_symm(cls, group, nbytes, device):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        t = symm.empty(nbytes, dtype=torch.uint8, device=device)
        hdl = symm.rendezvous(t, group)
        t.zero_()
        mc = int(hdl.multicast_ptr or 0)     # 0 where the fabric / driver offers no multicast object
        return cls(dist.get_world_size(group), dist.get_rank(group), list(hdl.buffer_ptrs), t,
                   multicast_ptr=mc, transport="symm", keepalive=(t, hdl))

    @classmethod
    def _create_ipc(cls, group, nbytes, device):
        import torch.distributed as dist
        lib = _lib.load()
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        with torch.cuda.device(device):
            ptr, handle = C.c_void_p(), C.create_string_buffer(64)
            _lib.check("b200attn_peer_alloc", lib.b200attn_peer_alloc(nbytes, C.byref(ptr), handle))
            handles = [None] * world
            dist.all_gather_object(handles, handle.raw, group=group)
            ptrs = []
            for r, h in enumerate(handles):
                if r == rank:
                    ptrs.append(ptr.value)
                    continue
                p = C.c_void_p()
                _lib.check("b200attn_peer_open", lib.b200attn_peer_open(C.create_string_buffer(h, 64), C.byref(p)))
                ptrs.append(p.value)
            local = torch.as_tensor(_RawCuda(ptr.value, nbytes), device=device)
        return cls(world, rank, ptrs, local, transport="ipc", keepalive=("ipc", ptr.value))

    @classmethod
    def local_ranks(cls, world: int, nbytes: int, device="cuda"):
        """`world` 'ranks' whose peers are plain allocations on ONE device — the protocol test bed: launch
        each rank's kernels on its own stream and they synchronise with each other exactly as ranks on
        different GPUs would (needs all the grids co-resident, which these small grids are)."""
        bufs = [torch.zeros(nbytes, dtype=torch.uint8, device=device) for _ in range(world)]
        ptrs = [b.data_ptr() for b in bufs]
        return [cls(world, r, ptrs, bufs[r], transport="local", keepalive=bufs) for r in range(world)]


class PeerAllReduce:
    """bf16 sum over the ranks of `group` through peer memory.

        ar = PeerAllReduce.create(group, max_numel=B * hidden)
        x = ar.input((B, hidden))        # the producer (o_proj) writes its partial sums here
        y = ar.all_reduce(x)             # one kernel launch on the current stream; y is local memory

    `all_reduce` accepts any bf16 CUDA tensor; one that is not the `input()` view is first copied
    into it (an extra device copy). Results of the one-/two-shot schedules are bit-identical on every
    rank: fp32 sum in rank order, one rounding to bf16.
    """

    def __init__(self, mem: PeerMemory, max_numel: int, timeout_s: float = 10.0, max_blocks: int = 0):
        if mem.world not in (1, 2, 4, 8):
            raise ValueError(f"world {mem.world}: 1, 2, 4 or 8 ranks")
        self.lib = _lib.load()
        self.mem = mem
        self.world, self.rank, self.max_numel = mem.world, mem.rank, max_numel
        self.off_in, self.off_scr, self.cap, self.total = region_layout(max_numel, mem.world)
        if mem.local.numel() < self.total:
            raise ValueError("allocation smaller than region_layout(max_numel, world)")
        self.push_slot_bytes = push_layout(max_numel, mem.world)
        self.off_push = self.off_scr + self.cap
        self.local = mem.local
        self.base_ptrs = mem.base_ptrs
        self.multicast_ptr = mem.multicast_ptr
        self.transport = mem.transport
        self.timeout_ns = int(timeout_s * 1e9)
        self.max_blocks = max_blocks
        self._in = _ptr_array([p + self.off_in for p in self.base_ptrs])
        self._scr = _ptr_array([p + self.off_scr for p in self.base_ptrs])
        self._flags = _ptr_array(self.base_ptrs)
        self._push = _ptr_array([p + self.off_push for p in self.base_ptrs]) if self.push_slot_bytes else None
        self._out_default = None
        self.last_algo = None
        if self.push_slot_bytes:   # "not arrived" sentinel: bf16 -0.0 in every half word of the receive buffers
            self.local[self.off_push: self.off_push + 3 * self.world * self.push_slot_bytes].view(torch.int16).fill_(-32768)

    @classmethod
    def create(cls, group=None, max_numel: int = 0, device=None, transport: str = "auto", **kw):
        import torch.distributed as dist
        grp = group if group is not None else dist.group.WORLD
        world = dist.get_world_size(grp)
        ar = cls(PeerMemory.create(group, region_layout(max_numel, world)[3], device, transport), max_numel, **kw)
        torch.cuda.synchronize()
        dist.barrier(grp)            # every rank's receive buffers hold the sentinel before anyone pushes
        return ar

    @classmethod
    def local_ranks(cls, world: int, max_numel: int, device="cuda", **kw):
        ars = [cls(m, max_numel, **kw)
               for m in PeerMemory.local_ranks(world, region_layout(max_numel, world)[3], device)]
        torch.cuda.synchronize()
        return ars

    # ---- use ----------------------------------------------------------------------------------
    @property
    def has_multicast(self) -> bool:
        return self.multicast_ptr != 0

    def input(self, shape) -> torch.Tensor:
        """bf16 view of this rank's input region (peers read it during `all_reduce`)."""
        n = 1
        for s in shape:
            n *= int(s)
        if n > self.max_numel:
            raise ValueError(f"{n} elements exceed max_numel {self.max_numel}")
        return self.local[self.off_in: self.off_in + 2 * n].view(torch.bfloat16).view(*shape)

    def output(self, shape) -> torch.Tensor:
        """bf16 view of this rank's output region (where MULTIMEM results land)."""
        n = 1
        for s in shape:
            n *= int(s)
        return self.local[self.off_scr: self.off_scr + 2 * n].view(torch.bfloat16).view(*shape)

    def error_words(self) -> list[int]:
        """[flag, block, peer, epoch] recorded by a kernel whose flag wait timed out (all 0 = none)."""
        return self.local[ERROR_WORDS_OFFSET: ERROR_WORDS_OFFSET + 16].view(torch.int32).tolist()

    def all_reduce(self, x: torch.Tensor, out: torch.Tensor | None = None, algo: int = _lib.AR_AUTO) -> torch.Tensor:
        if not x.is_cuda or x.dtype != torch.bfloat16 or not x.is_contiguous():
            raise ValueError("all_reduce expects a contiguous bf16 CUDA tensor (there is no CPU path)")
        n = x.numel()
        if n % 8 or n > self.max_numel:
            raise ValueError(f"numel {n}: multiple of 8, at most {self.max_numel}")
        if algo == _lib.AR_AUTO:
            algo = choose_algo(2 * n, self.world, self.has_multicast, self.push_slot_bytes)
        in_ptrs = self._in
        if algo in (_lib.AR_PUSH, _lib.AR_PUSH2):     # read only the local input: any device memory, no staging copy
            if x.data_ptr() != self.base_ptrs[self.rank] + self.off_in:
                ptrs = [p + self.off_in for p in self.base_ptrs]
                ptrs[self.rank] = x.data_ptr()
                in_ptrs = _ptr_array(ptrs)
        elif x.data_ptr() != self.base_ptrs[self.rank] + self.off_in:
            self.input(x.shape).copy_(x)
        mc_in = mc_out = None
        if algo == _lib.AR_MULTIMEM:
            if not self.has_multicast:
                raise PeerSetupError("MULTIMEM requested but the allocation has no multicast mapping")
            mc_in, mc_out = self.multicast_ptr + self.off_in, self.multicast_ptr + self.off_scr
            res = self.output(x.shape)
        else:
            if out is None:
                if self._out_default is None or self._out_default.numel() < n:
                    self._out_default = torch.empty(self.max_numel, dtype=torch.bfloat16, device=x.device)
                out = self._out_default[:n].view(x.shape)
            res = out
        stream = torch.cuda.current_stream(x.device).cuda_stream
        _lib.check("b200attn_peer_allreduce", self.lib.b200attn_peer_allreduce(
            in_ptrs, self._scr, self._flags, self._push, self.push_slot_bytes, res.data_ptr(), mc_in, mc_out, n,
            self.world, self.rank, algo,
            self.max_blocks, self.timeout_ns, stream))
        self.last_algo = algo
        if algo == _lib.AR_MULTIMEM and out is not None:
            out.copy_(res)
            return out
        return res


class PeerExchange:
    """q / output row exchange inside a KV-head replica group through peer memory (the device side of
    `tp.ReplicaSplit`, replacing its two NCCL all_to_all calls): `b200attn_peer_exchange`.

        ex = PeerExchange.create(group, batch, hq_rank, d)
        ex.q[...]  = this rank's query heads [batch, hq_rank, d]   (the producer writes here)
        ex.gather_q()        -> ex.q_grp   [per, R * hq_rank, d]: all heads of the group, my sequences
        decode(ex.q_grp[:n]) -> ex.out_grp [per, R * hq_rank, d]
        ex.scatter_out()     -> ex.out     [batch, hq_rank, d]: my heads of every sequence
    """

    def __init__(self, mem: PeerMemory, batch: int, hq: int, d: int, timeout_s: float = 10.0):
        self.lib = _lib.load()
        self.mem, self.R, self.index = mem, mem.world, mem.rank
        self.batch, self.hq, self.d = batch, hq, d
        self.row_bytes = hq * d * 2
        self.per = -(-batch // self.R)
        self.b0 = min(batch, self.index * self.per)
        self.b1 = min(batch, self.b0 + self.per)
        off_q, off_out, total = self.layout(batch, hq, d)
        if mem.local.numel() < total:
            raise ValueError("allocation smaller than PeerExchange.layout()")
        nb = batch * self.row_bytes
        self.q = mem.local[off_q: off_q + nb].view(torch.bfloat16).view(batch, hq, d)
        self.out = mem.local[off_out: off_out + nb].view(torch.bfloat16).view(batch, hq, d)
        dev = mem.local.device
        self.q_grp = torch.zeros(self.per, self.R * hq, d, dtype=torch.bfloat16, device=dev)
        self.out_grp = torch.zeros(self.per, self.R * hq, d, dtype=torch.bfloat16, device=dev)
        self._flags = _ptr_array(mem.base_ptrs)
        self._q = _ptr_array([p + off_q for p in mem.base_ptrs])
        self._out = _ptr_array([p + off_out for p in mem.base_ptrs])
        self.timeout_ns = int(timeout_s * 1e9)

    @staticmethod
    def layout(batch: int, hq: int, d: int) -> tuple[int, int, int]:
        """(q offset, out offset, total bytes) of one rank's allocation: [flags | q | out]."""
        nb = -(-(batch * hq * d * 2) // _ALIGN) * _ALIGN
        return _lib.PEER_FLAG_BYTES, _lib.PEER_FLAG_BYTES + nb, _lib.PEER_FLAG_BYTES + 2 * nb

    @classmethod
    def create(cls, group, batch: int, hq: int, d: int, device=None, transport: str = "auto", **kw):
        return cls(PeerMemory.create(group, cls.layout(batch, hq, d)[2], device, transport), batch, hq, d, **kw)

    @classmethod
    def local_ranks(cls, R: int, batch: int, hq: int, d: int, device="cuda", **kw):
        return [cls(m, batch, hq, d, **kw) for m in PeerMemory.local_ranks(R, cls.layout(batch, hq, d)[2], device)]

    def split(self, batch: int | None = None) -> tuple[int, int, int]:
        """(rows per rank incl. padding, first local row, one past the last) for a live batch <= the one sized for."""
        n = self.batch if batch is None else int(batch)
        if n > self.batch:
            raise ValueError(f"exchange was sized for batch {self.batch}, got {n}")
        per = -(-n // self.R)
        b0 = min(n, self.index * per)
        return per, b0, min(n, b0 + per)

    def _run(self, ptrs, local, scatter: int, batch: int | None):
        per, b0, b1 = self.split(batch)
        stream = torch.cuda.current_stream(local.device).cuda_stream
        _lib.check("b200attn_peer_exchange", self.lib.b200attn_peer_exchange(
            ptrs, self._flags, local.data_ptr(), self.R, self.index, b0, b1 - b0, per,
            self.row_bytes, scatter, self.timeout_ns, stream))

    def gather_q(self, batch: int | None = None) -> torch.Tensor:
        """All ranks of the group call it with the same live batch (rows [0, batch) of `q` are exchanged)."""
        self._run(self._q, self.q_grp, 0, batch)
        return self.q_grp

    def scatter_out(self, batch: int | None = None) -> torch.Tensor:
        self._run(self._out, self.out_grp, 1, batch)
        return self.out
