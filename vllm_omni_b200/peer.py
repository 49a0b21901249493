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
import os

import torch

from . import _lib

ALGO_NAMES = {_lib.AR_AUTO: "auto", _lib.AR_ONESHOT: "oneshot", _lib.AR_TWOSHOT: "twoshot",
              _lib.AR_MULTIMEM: "multimem"}
_ALIGN = 256


class PeerSetupError(RuntimeError):
    """Peer-mapped memory could not be established between the ranks."""


def region_layout(max_numel: int) -> tuple[int, int, int, int]:
    """(input offset, scratch/out offset, region capacity in bytes, total bytes) of one rank's allocation
    for bf16 tensors of up to `max_numel` elements. The scratch region doubles as the multimem output
    region, so it has the full capacity."""
    cap = -(-(max_numel * 2 + 16) // _ALIGN) * _ALIGN
    off_in = _lib.PEER_FLAG_BYTES
    off_scr = off_in + cap
    return off_in, off_scr, cap, off_scr + cap


def choose_algo(nbytes: int, world: int, has_multicast: bool) -> int:
    """The schedule `AUTO` resolves to (host arithmetic in the C library, no GPU needed)."""
    return int(_lib.load().b200attn_peer_allreduce_algo(nbytes, world, int(bool(has_multicast))))


class _RawCuda:
    """`__cuda_array_interface__` view of raw device memory (lets torch wrap a cudaIpc pointer)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False),
                                         "version": 3, "strides": None}


def _ptr_array(ptrs):
    return (C.c_void_p * len(ptrs))(*[C.c_void_p(int(p)) for p in ptrs])


class PeerAllReduce:
    """bf16 sum over the ranks of `group` through peer memory.

        ar = PeerAllReduce.create(group, max_numel=B * hidden)
        x = ar.input((B, hidden))        # the producer (o_proj) writes its partial sums here
        y = ar.all_reduce(x)             # one kernel launch on the current stream; y is local memory

    `all_reduce` accepts any bf16 CUDA tensor; one that is not the `input()` view is first copied
    into it (an extra device copy). Results of the one-/two-shot schedules are bit-identical on every
    rank: fp32 sum in rank order, one rounding to bf16.
    """

    def __init__(self, world: int, rank: int, base_ptrs, max_numel: int, local: torch.Tensor,
                 multicast_ptr: int = 0, transport: str = "local", keepalive=None,
                 timeout_s: float = 10.0, max_blocks: int = 0):
        if world not in (1, 2, 4, 8):
            raise ValueError(f"world {world}: 1, 2, 4 or 8 ranks")
        self.lib = _lib.load()
        self.world, self.rank, self.max_numel = world, rank, max_numel
        self.off_in, self.off_scr, self.cap, self.total = region_layout(max_numel)
        if local.numel() * local.element_size() < self.total:
            raise ValueError("allocation smaller than region_layout(max_numel)")
        self.local = local.view(torch.uint8) if local.dtype != torch.uint8 else local   # this rank's allocation
        self.base_ptrs = [int(p) for p in base_ptrs]
        self.multicast_ptr = int(multicast_ptr or 0)
        self.transport = transport
        self.timeout_ns = int(timeout_s * 1e9)
        self.max_blocks = max_blocks
        self._keep = keepalive
        self._in = _ptr_array([p + self.off_in for p in self.base_ptrs])
        self._scr = _ptr_array([p + self.off_scr for p in self.base_ptrs])
        self._flags = _ptr_array(self.base_ptrs)
        self._out_default = None
        self.last_algo = None

    # ---- construction -------------------------------------------------------------------------
    @classmethod
    def create(cls, group=None, max_numel: int = 0, device=None, transport: str = "auto", **kw):
        """Collective over `group` (default: WORLD). transport: 'symm' (torch symmetric memory; gives a
        multicast mapping where the fabric has one), 'ipc' (cudaMalloc + cudaIpc handles) or 'auto'
        (symm, then ipc)."""
        import torch.distributed as dist
        group = group if group is not None else dist.group.WORLD
        device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        errors = []
        for tr in (("symm", "ipc") if transport == "auto" else (transport,)):
            try:
                ar = (cls._create_symm if tr == "symm" else cls._create_ipc)(group, max_numel, device, **kw)
            except Exception as ex:  # noqa: BLE001 - every rank must take the same branch: vote below
                ar, err = None, f"{tr}: {type(ex).__name__}: {ex}"
            else:
                err = None
            votes = [None] * dist.get_world_size(group)
            dist.all_gather_object(votes, err, group=group)
            if all(v is None for v in votes):
                torch.cuda.synchronize(device)
                dist.barrier(group)          # every rank's flag region is zeroed before anyone launches
                return ar
            errors.append("; ".join(sorted({v for v in votes if v})))
        raise PeerSetupError(" | ".join(errors))

    @classmethod
    def _create_symm(cls, group, max_numel, device, **kw):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        total = region_layout(max_numel)[3]
        t = symm.empty(total, dtype=torch.uint8, device=device)
        hdl = symm.rendezvous(t, group)
        t.zero_()
        mc = int(hdl.multicast_ptr or 0)     # 0 where the fabric / driver offers no multicast object
        return cls(dist.get_world_size(group), dist.get_rank(group), list(hdl.buffer_ptrs), max_numel, t,
                   multicast_ptr=mc, transport="symm", keepalive=(t, hdl), **kw)

    @classmethod
    def _create_ipc(cls, group, max_numel, device, **kw):
        import torch.distributed as dist
        lib = _lib.load()
        total = region_layout(max_numel)[3]
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        with torch.cuda.device(device):
            ptr, handle = C.c_void_p(), C.create_string_buffer(64)
            _lib.check("b200attn_peer_alloc", lib.b200attn_peer_alloc(total, C.byref(ptr), handle))
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
            local = torch.as_tensor(_RawCuda(ptr.value, total), device=device)
        return cls(world, rank, ptrs, max_numel, local, transport="ipc", keepalive=("ipc", ptr.value), **kw)

    @classmethod
    def local_ranks(cls, world: int, max_numel: int, device="cuda", **kw):
        """`world` instances whose 'peers' are plain allocations on ONE device — the protocol test bed:
        launch each on its own stream and the kernels synchronise with each other exactly as ranks on
        different GPUs would (needs all `world` grids co-resident, which small grids are)."""
        total = region_layout(max_numel)[3]
        bufs = [torch.zeros(total, dtype=torch.uint8, device=device) for _ in range(world)]
        ptrs = [b.data_ptr() for b in bufs]
        return [cls(world, r, ptrs, max_numel, bufs[r], transport="local", keepalive=bufs, **kw)
                for r in range(world)]

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
        off = (128 * 8 + 128) * 4
        return self.local[off: off + 16].view(torch.int32).tolist()

    def all_reduce(self, x: torch.Tensor, out: torch.Tensor | None = None, algo: int = _lib.AR_AUTO) -> torch.Tensor:
        if not x.is_cuda or x.dtype != torch.bfloat16 or not x.is_contiguous():
            raise ValueError("all_reduce expects a contiguous bf16 CUDA tensor (there is no CPU path)")
        n = x.numel()
        if n % 8 or n > self.max_numel:
            raise ValueError(f"numel {n}: multiple of 8, at most {self.max_numel}")
        if x.data_ptr() != self.base_ptrs[self.rank] + self.off_in:
            self.input(x.shape).copy_(x)
        if algo == _lib.AR_AUTO:
            algo = choose_algo(2 * n, self.world, self.has_multicast)
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
            self._in, self._scr, self._flags, res.data_ptr(), mc_in, mc_out, n, self.world, self.rank, algo,
            self.max_blocks, self.timeout_ns, stream))
        self.last_algo = algo
        if algo == _lib.AR_MULTIMEM and out is not None:
            out.copy_(res)
            return out
        return res
