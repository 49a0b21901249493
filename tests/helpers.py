"""Test-side construction of small paged-attention cases (CPU tensors; the oracle fills the cache)."""
from __future__ import annotations

from dataclasses import dataclass

import torch

from oracle import oracle
from vllm_omni_b200 import layout
from vllm_omni_b200.workloads import default_scales, make_block_table, slots_of_positions


@dataclass
class Case:
    q: torch.Tensor            # [B,Hq,d] bf16
    hist_k: torch.Tensor       # [sum S, Hkv, d] bf16, token-major history (all positions)
    hist_v: torch.Tensor
    slots: torch.Tensor        # [sum S] int64
    block_table: torch.Tensor  # [B,max_blocks] int32
    seq_lens: torch.Tensor     # [B] int32
    num_blocks: int
    block_size: int
    fp8: bool
    k_scale: torch.Tensor | None
    v_scale: torch.Tensor | None
    scale: float
    kv_cache: torch.Tensor     # CPU cache filled by the oracle, logical [2,nb,bs,Hkv,d]

    @property
    def max_seq_len(self) -> int:
        return int(self.seq_lens.max()) if self.seq_lens.numel() else 0

    @property
    def kv_cache_dtype(self) -> str:
        return "fp8_e4m3" if self.fp8 else "auto"


def make_case(seq_lens, hq=8, hkv=2, d=128, bs=16, seed=0, fp8=False, per_head_scale=True,
              contiguous=False, q_std=1.0, kv_std=1.0, garbage_fill=False) -> Case:
    g = torch.Generator().manual_seed(seed)
    lens = torch.as_tensor(seq_lens, dtype=torch.int32)
    B = lens.numel()
    bt, nb = make_block_table(lens, bs, spare_blocks=3, contiguous=contiguous)
    total = int(lens.sum())
    q = (torch.randn(B, hq, d, generator=g) * q_std).to(torch.bfloat16)
    hk = (torch.randn(total, hkv, d, generator=g) * kv_std).to(torch.bfloat16)
    hv = (torch.randn(total, hkv, d, generator=g) * kv_std).to(torch.bfloat16)
    l64 = lens.to(torch.int64)
    seq_idx = torch.repeat_interleave(torch.arange(B), l64)
    starts = torch.cumsum(l64, 0) - l64
    pos = torch.arange(total) - torch.repeat_interleave(starts, l64)
    slots = slots_of_positions(bt, seq_idx, pos, bs) if total else torch.zeros(0, dtype=torch.int64)
    ks = vs = None
    if fp8:
        ks, vs = default_scales(hkv)
        if not per_head_scale:
            ks, vs = ks[:1].clone(), vs[:1].clone()
    cache = layout.alloc_kv_cache(nb, bs, hkv, d, "fp8_e4m3" if fp8 else "auto", "cpu")
    if garbage_fill:
        # unused slots hold non-finite bit patterns: masked tokens must not leak into the output
        if fp8:
            cache.fill_(0x7F)
        else:
            cache.view(torch.int16).fill_(0x7FC0)  # bf16 NaN
    oracle.ref_reshape_and_cache(hk, hv, cache[0], cache[1], slots, fp8, ks, vs)
    return Case(q, hk, hv, slots, bt, lens, nb, bs, fp8, ks, vs, d ** -0.5, cache)


def to_dev(t, dev="cuda"):
    return None if t is None else t.to(dev)


def cache_to_dev(cache_cpu: torch.Tensor, dev="cuda") -> torch.Tensor:
    """Copy a CPU cache to the device keeping the backend's physical layout."""
    nb, bs, hkv, d = cache_cpu.shape[1:]
    kind = "fp8_e4m3" if cache_cpu.dtype == torch.uint8 else "auto"
    out = layout.alloc_kv_cache(nb, bs, hkv, d, kind, dev)
    out.copy_(cache_cpu)
    return out


def bf16_half_ulp(ref: torch.Tensor) -> torch.Tensor:
    """Half a bf16 ulp at |ref| (the unavoidable rounding of a bf16 output)."""
    a = ref.abs().to(torch.float32).clamp_min(2.0 ** -126)
    e = torch.floor(torch.log2(a))
    return torch.pow(2.0, e - 8)


def excess_rel_err(x: torch.Tensor, ref: torch.Tensor) -> float:
    """max(|x - ref| - half_ulp_bf16(ref)) / max|ref|: arithmetic error beyond the mandatory bf16
    rounding of the output, relative to the output scale. The north_star's 2e-3 bound is applied
    to this number (a bf16 output cannot meet 2e-3 element-wise: its own rounding is up to 3.9e-3)."""
    x = x.detach().to(torch.float32).cpu()
    ref = ref.detach().to(torch.float32).cpu()
    err = ((x - ref).abs() - bf16_half_ulp(ref)).clamp_min(0)
    return float(err.max()) / float(ref.abs().max().clamp_min(1e-20))


def row_excess_rel_err(x: torch.Tensor, ref: torch.Tensor) -> float:
    """The same excess error, but every output row ([..., d]: one (sequence, head)) is divided by ITS OWN max|ref|,
    so a row of small magnitude cannot hide behind a large one (max over rows)."""
    x = x.detach().to(torch.float32).cpu()
    ref = ref.detach().to(torch.float32).cpu()
    err = ((x - ref).abs() - bf16_half_ulp(ref)).clamp_min(0)
    scale = ref.abs().amax(dim=-1, keepdim=True)
    ok = scale > 0
    return float((err / scale.clamp_min(1e-30))[ok.expand_as(err)].max()) if bool(ok.any()) else 0.0


def rel_l2(x: torch.Tensor, ref: torch.Tensor) -> float:
    x = x.detach().to(torch.float32).cpu()
    ref = ref.detach().to(torch.float32).cpu()
    return float((x - ref).norm() / ref.norm().clamp_min(1e-20))


def kernel_rel_l2(x: torch.Tensor, ref: torch.Tensor) -> float:
    """Relative L2 error attributable to the kernel's arithmetic: the total rel-L2 of the bf16
    output with the rel-L2 of the mandatory output rounding, rel_l2(bf16(ref), ref) ~ 1.65e-3,
    removed in quadrature. A bf16 tensor cannot be closer than that to an fp32 reference, so the
    2e-3 bound of the north_star is applied to what the kernel adds on top (in practice the bf16
    rounding of P before the PV tensor-core product, ~1.3e-3)."""
    ref32 = ref.detach().to(torch.float32).cpu()
    total = rel_l2(x, ref32)
    floor = rel_l2(ref32.to(torch.bfloat16), ref32)
    return max(0.0, total * total - floor * floor) ** 0.5
