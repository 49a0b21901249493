"""Tensor-level host wrappers over the C ABI (include/b200attn.h).

PyTorch is used for device memory and streams only; every op here marshals raw pointers,
element strides and torch's current CUDA stream into libb200attn.so. There is no fallback:
CPU tensors are rejected and a missing library raises `B200AttnLibraryError`.
"""
from __future__ import annotations

import torch

from . import _lib
from ._lib import (DECODE_EARLY_KV_READ, DECODE_EARLY_Q_READ, DECODE_PLANNED, KV_BF16, KV_FP8_E4M3, VARIANT_AUTO, VARIANT_CPASYNC,  # noqa: F401
                   VARIANT_SK2, VARIANT_SK3, VARIANT_STREAM, VARIANT_STREAMK, VARIANT_TMA, check)

WORKSPACE_MIN_BYTES = 65536  # control area (arrival / claim counters), see b200attn_workspace_init

_FP8_NAMES = ("fp8", "fp8_e4m3")


def kv_dtype_code(kv_cache_dtype: str | int) -> int:
    """vLLM cache dtype string ("auto", "bfloat16", "fp8", "fp8_e4m3") -> ABI code."""
    if isinstance(kv_cache_dtype, int):
        return kv_cache_dtype
    if kv_cache_dtype in _FP8_NAMES:
        return KV_FP8_E4M3
    if kv_cache_dtype in ("auto", "bfloat16", "bf16"):
        return KV_BF16
    raise ValueError(f"kv_cache_dtype {kv_cache_dtype!r} is not supported (bf16 or fp8_e4m3)")


def split_kv_cache(kv_cache: torch.Tensor) -> tuple[torch.Tensor, torch.Tensor]:
    """Logical [2, nb, bs, Hkv, d] (or [nb, 2, bs, Hkv, d]) -> (k_cache, v_cache) views [nb, bs, Hkv, d].

    Same normalisation rule as the reference's reader
    (vllm_omni/distributed/omni_connectors/utils/kv_utils.py:417-450).
    """
    if kv_cache.ndim != 5:
        raise ValueError(f"kv_cache must be 5-D, got {tuple(kv_cache.shape)}")
    if kv_cache.shape[0] == 2:
        return kv_cache[0], kv_cache[1]
    if kv_cache.shape[1] == 2:
        return kv_cache[:, 0], kv_cache[:, 1]
    raise ValueError(f"kv_cache has no K/V pair at dim 0 or 1: {tuple(kv_cache.shape)}")


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _need_cuda(*ts: torch.Tensor | None) -> None:
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError("b200attn ops run on CUDA tensors only (no CPU fallback)")


def _cache_strides(k_cache: torch.Tensor, v_cache: torch.Tensor) -> tuple[int, int, int]:
    if k_cache.ndim != 4 or k_cache.shape != v_cache.shape or k_cache.stride() != v_cache.stride():
        raise ValueError("k_cache and v_cache must be [nb, bs, Hkv, d] views with equal strides")
    if k_cache.stride(3) != 1:
        raise ValueError("head_size must be the contiguous cache dimension")
    return k_cache.stride(0), k_cache.stride(1), k_cache.stride(2)


def _scales(k_scale, v_scale, fp8: bool):
    if not fp8:
        return None, None, 1
    if k_scale is None or v_scale is None:
        raise ValueError("fp8 KV cache needs k_scale and v_scale")
    if k_scale.dtype != torch.float32 or v_scale.dtype != torch.float32:
        raise ValueError("scales must be float32 device tensors")
    if k_scale.numel() != v_scale.numel():
        raise ValueError("k_scale and v_scale must have the same number of elements")
    _need_cuda(k_scale, v_scale)
    return k_scale.data_ptr(), v_scale.data_ptr(), k_scale.numel()


def num_sms() -> int:
    return _lib.load().b200attn_num_sms()


def reshape_and_cache(key: torch.Tensor, value: torch.Tensor, k_cache: torch.Tensor,
                      v_cache: torch.Tensor, slot_mapping: torch.Tensor,
                      kv_cache_dtype: str | int = "auto", k_scale: torch.Tensor | None = None,
                      v_scale: torch.Tensor | None = None) -> None:
    """KV-cache write; same call shape as vLLM `reshape_and_cache_flash`
    (vllm/_custom_ops.py:2591-2610). Token count comes from `slot_mapping` (key/value may be padded)."""
    lib = _lib.load()
    _need_cuda(key, value, k_cache, v_cache, slot_mapping)
    code = kv_dtype_code(kv_cache_dtype)
    fp8 = code == KV_FP8_E4M3
    if key.dtype != torch.bfloat16 or value.dtype != torch.bfloat16:
        raise ValueError("key/value must be bfloat16")
    if key.ndim != 3 or value.ndim != 3 or key.stride(2) != 1 or value.stride(2) != 1:
        raise ValueError("key/value must be [T, Hkv, d] with contiguous d")
    if slot_mapping.dtype != torch.int64 or not slot_mapping.is_contiguous():
        raise ValueError("slot_mapping must be contiguous int64")
    n = slot_mapping.numel()
    if n > key.shape[0] or n > value.shape[0]:
        raise ValueError("slot_mapping longer than key/value")
    want = torch.uint8 if fp8 else torch.bfloat16
    if fp8 and k_cache.dtype in (torch.float8_e4m3fn,):
        k_cache, v_cache = k_cache.view(torch.uint8), v_cache.view(torch.uint8)
    if k_cache.dtype != want or v_cache.dtype != want:
        raise ValueError(f"cache dtype {k_cache.dtype} does not match kv_cache_dtype")
    sb, ss, sh = _cache_strides(k_cache, v_cache)
    nb, bs, hkv, d = k_cache.shape
    if key.shape[1] != hkv or key.shape[2] != d or value.shape[1:] != key.shape[1:]:
        raise ValueError("key/value head layout does not match the cache")
    ks, vs, sn = _scales(k_scale, v_scale, fp8)
    check("b200attn_reshape_and_cache", lib.b200attn_reshape_and_cache(
        key.data_ptr(), value.data_ptr(), key.stride(0), key.stride(1), value.stride(0),
        value.stride(1), k_cache.data_ptr(), v_cache.data_ptr(), sb, ss, sh,
        slot_mapping.data_ptr(), n, hkv, d, bs, nb, code, ks, vs, sn, _stream()))


def rope_and_cache(positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor, value: torch.Tensor,
                   cos_sin_cache: torch.Tensor, is_neox: bool, k_cache: torch.Tensor, v_cache: torch.Tensor,
                   slot_mapping: torch.Tensor, kv_cache_dtype: str | int = "auto",
                   k_scale: torch.Tensor | None = None, v_scale: torch.Tensor | None = None,
                   rot_dim: int | None = None, mrope_section: list[int] | None = None,
                   mrope_interleaved: bool = False) -> None:
    """Fused rotary embedding (query and key rotated IN PLACE) + KV-cache write of the rotated key and
    the value; one launch for what vLLM does as rotary_emb + reshape_and_cache_flash
    (hook: AttentionImpl.do_rope_and_kv_cache_update, vllm/v1/attention/backend.py:815-840).
    query [T, Hq, d] / key, value [T, Hkv, d] bf16 (token/head strides free), positions int64 [T] or [3, T]
    (M-RoPE), cos_sin_cache bf16 [max_pos, rot_dim] = cos | sin, slot_mapping int64 [<= T]."""
    import ctypes as C
    lib = _lib.load()
    _need_cuda(positions, query, key, value, cos_sin_cache, k_cache, v_cache, slot_mapping)
    code = kv_dtype_code(kv_cache_dtype)
    fp8 = code == KV_FP8_E4M3
    for name, x in (("query", query), ("key", key), ("value", value)):
        if x.dtype != torch.bfloat16 or x.ndim != 3 or x.stride(2) != 1:
            raise ValueError(f"{name} must be bfloat16 [T, H, d] with contiguous d")
    if query.stride(1) != query.shape[2]:
        raise ValueError("query heads must be contiguous")
    if cos_sin_cache.dtype != torch.bfloat16 or cos_sin_cache.ndim != 2 or cos_sin_cache.stride(1) != 1:
        raise ValueError("cos_sin_cache must be bfloat16 [max_pos, rot_dim] (cast it once at load time)")
    if positions.dtype != torch.int64 or positions.ndim not in (1, 2) or positions.stride(-1) != 1:
        raise ValueError("positions must be int64 [T] or [3, T]")
    if slot_mapping.dtype != torch.int64 or not slot_mapping.is_contiguous():
        raise ValueError("slot_mapping must be contiguous int64")
    T = positions.shape[-1]
    if query.shape[0] < T or key.shape[0] < T or value.shape[0] < T or slot_mapping.numel() > T:
        raise ValueError("token counts of positions / query / key / value / slot_mapping do not agree")
    rot = int(rot_dim) if rot_dim is not None else cos_sin_cache.shape[1]
    if fp8 and k_cache.dtype == torch.float8_e4m3fn:
        k_cache, v_cache = k_cache.view(torch.uint8), v_cache.view(torch.uint8)
    want = torch.uint8 if fp8 else torch.bfloat16
    if k_cache.dtype != want or v_cache.dtype != want:
        raise ValueError(f"cache dtype {k_cache.dtype} does not match kv_cache_dtype")
    sb, ss, sh = _cache_strides(k_cache, v_cache)
    nb, bs, hkv, d = k_cache.shape
    if key.shape[1] != hkv or key.shape[2] != d or value.shape[1:] != key.shape[1:] or query.shape[2] != d:
        raise ValueError("query/key/value head layout does not match the cache")
    rows = 1 if positions.ndim == 1 else positions.shape[0]
    sec = None
    if rows != 1:
        if mrope_section is None or len(mrope_section) != 3:
            raise ValueError("[3, T] positions need mrope_section (3 ints)")
        sec = (C.c_int32 * 3)(*[int(x) for x in mrope_section])
    ks, vs, sn = _scales(k_scale, v_scale, fp8)
    check("b200attn_rope_and_cache", lib.b200attn_rope_and_cache(
        query.data_ptr(), query.stride(0), query.shape[1], key.data_ptr(), value.data_ptr(), key.stride(0),
        key.stride(1), value.stride(0), value.stride(1), positions.data_ptr(), rows,
        positions.stride(0) if rows != 1 else 0, cos_sin_cache.data_ptr(), cos_sin_cache.stride(0),
        cos_sin_cache.shape[0], rot, 1 if is_neox else 0, sec, 1 if mrope_interleaved else 0,
        k_cache.data_ptr(), v_cache.data_ptr(), sb, ss, sh, slot_mapping.data_ptr(), T, slot_mapping.numel(),
        hkv, d, bs, nb, code, ks, vs, sn, _stream()))


def decode_num_splits(batch: int, num_q_heads: int, num_kv_heads: int, max_seq_len: int,
                      sms: int | None = None) -> int:
    lib = _lib.load()
    return lib.b200attn_decode_num_splits(batch, num_q_heads, num_kv_heads, max_seq_len,
                                          sms if sms is not None else num_sms())


def decode_workspace_bytes(max_batch: int, num_q_heads: int, head_size: int, max_splits: int) -> int:
    return _lib.load().b200attn_decode_workspace_bytes(max_batch, num_q_heads, head_size, max_splits)


class DecodeWorkspace:
    """Caller-owned split-K scratch (control area, partial outputs + LSE, the step's stream-K plan), grown
    on demand outside graphs. The control area (first 16 KiB) is zeroed once per allocation through
    `b200attn_workspace_init`; the library keeps it zero between launches."""

    def __init__(self, device: torch.device | str = "cuda", nbytes: int = 0):
        self.device = torch.device(device)
        self._alloc(max(nbytes, WORKSPACE_MIN_BYTES))

    def _alloc(self, nbytes: int) -> None:
        self.buf = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
        if self.buf.is_cuda:
            with torch.cuda.device(self.device):
                check("b200attn_workspace_init", _lib.load().b200attn_workspace_init(
                    self.buf.data_ptr(), self.buf.numel(), torch.cuda.current_stream(self.device).cuda_stream))

    def ensure(self, nbytes: int) -> None:
        if self.buf.numel() < nbytes:
            if torch.cuda.is_current_stream_capturing():
                raise RuntimeError("decode workspace too small and cannot grow during graph capture")
            self._alloc(nbytes)

    @property
    def ptr(self) -> int:
        return self.buf.data_ptr()

    @property
    def nbytes(self) -> int:
        return self.buf.numel()


def _decode_args(q, k_cache, v_cache, out, lse, block_table, seq_lens, max_seq_len, softmax_scale,
                 kv_cache_dtype, k_scale, v_scale, num_splits, variant, workspace, flags=0):
    _need_cuda(q, k_cache, v_cache, out, lse, block_table, seq_lens)
    code = kv_dtype_code(kv_cache_dtype)
    fp8 = code == KV_FP8_E4M3
    if q.dtype != torch.bfloat16 or out.dtype != torch.bfloat16:
        raise ValueError("q/out must be bfloat16")
    if q.ndim != 3 or q.stride(2) != 1 or q.stride(1) != q.shape[2]:
        raise ValueError("q must be [B, Hq, d] with contiguous heads")
    if out.shape != q.shape or out.stride(2) != 1 or out.stride(1) != out.shape[2]:
        raise ValueError("out must be [B, Hq, d] with contiguous heads")
    if fp8 and k_cache.dtype == torch.float8_e4m3fn:
        k_cache, v_cache = k_cache.view(torch.uint8), v_cache.view(torch.uint8)
    want = torch.uint8 if fp8 else torch.bfloat16
    if k_cache.dtype != want or v_cache.dtype != want:
        raise ValueError(f"cache dtype {k_cache.dtype} does not match kv_cache_dtype")
    sb, ss, sh = _cache_strides(k_cache, v_cache)
    nb, bs, hkv, d = k_cache.shape
    batch, hq, dq = q.shape
    if dq != d:
        raise ValueError("head_size mismatch between q and cache")
    if block_table.dtype != torch.int32 or block_table.ndim != 2 or block_table.stride(1) != 1:
        raise ValueError("block_table must be int32 [B, max_blocks] with contiguous rows")
    if seq_lens.dtype != torch.int32 or not seq_lens.is_contiguous():
        raise ValueError("seq_lens must be contiguous int32")
    if block_table.shape[0] < batch or seq_lens.numel() < batch:
        raise ValueError("block_table/seq_lens shorter than the batch")
    if lse is not None and (lse.dtype != torch.float32 or not lse.is_contiguous()
                            or lse.numel() < batch * hq):
        raise ValueError("lse must be contiguous float32 [B, Hq]")
    ks, vs, sn = _scales(k_scale, v_scale, fp8)
    ws_ptr, ws_bytes = (workspace.ptr, workspace.nbytes) if workspace is not None else (None, 0)
    return (q.data_ptr(), q.stride(0), k_cache.data_ptr(), v_cache.data_ptr(), sb, ss, sh, nb,
            out.data_ptr(), out.stride(0), lse.data_ptr() if lse is not None else None,
            block_table.data_ptr(), block_table.stride(0), seq_lens.data_ptr(), batch,
            int(max_seq_len), hq, hkv, d, bs, float(softmax_scale), code, ks, vs, sn,
            int(num_splits), int(variant), int(flags), ws_ptr, ws_bytes, _stream())


def decode_plan(seq_lens: torch.Tensor, block_table: torch.Tensor, batch: int, num_q_heads: int,
                num_kv_heads: int, block_size: int, workspace: DecodeWorkspace) -> None:
    """Stream-K plan of one decode step (one small launch per STEP, not per layer): every
    `paged_decode(..., planned=True)` of the step must be given the same seq_lens / block_table /
    batch / head counts / block size and this workspace."""
    lib = _lib.load()
    _need_cuda(seq_lens, block_table)
    if block_table.dtype != torch.int32 or block_table.ndim != 2 or block_table.stride(1) != 1:
        raise ValueError("block_table must be int32 [B, max_blocks] with contiguous rows")
    if seq_lens.dtype != torch.int32 or not seq_lens.is_contiguous():
        raise ValueError("seq_lens must be contiguous int32")
    if block_table.shape[0] < batch or seq_lens.numel() < batch:
        raise ValueError("block_table/seq_lens shorter than the batch")
    workspace.ensure(decode_workspace_bytes(batch, num_q_heads, 128, 1))
    check("b200attn_decode_plan", lib.b200attn_decode_plan(
        seq_lens.data_ptr(), int(batch), int(num_q_heads), int(num_kv_heads), block_table.data_ptr(),
        block_table.stride(0), int(block_size), workspace.ptr, workspace.nbytes, _stream()))


def debug_sk3_plan(workspace: DecodeWorkspace, batch: int, num_q_heads: int) -> dict:
    """The plan `decode_plan` left in `workspace`, parsed (tests / timeline scripts; synchronises)."""
    lib = _lib.load()
    off = int(lib.b200attn_debug_plan_offset(int(batch), int(num_q_heads)))
    words = workspace.buf[off:].view(torch.int32)
    hdr = words[:16].cpu().tolist()
    total, groups, T, nt, b, hs_count, _, magic, tpc = hdr[:9]
    if magic != 0x534B3350:
        raise RuntimeError("no cooperative stream-K plan in this workspace")
    ntmax = groups * 12
    o_desc = 16 + ((ntmax + 1 + 3) & ~3)
    n_words = o_desc + 44 * ntmax + batch + 1   # (32 block ids per warp sit between the warp descriptors and the prefix)
    w = words[:n_words].cpu()
    return {"total": total, "groups": groups, "T": T, "nt": nt, "batch": b, "hs_count": hs_count, "tpc": tpc,
            "tstart": w[16:16 + ntmax + 1].tolist(),
            "tdesc": w[o_desc:o_desc + 4 * ntmax].view(ntmax, 4).tolist(),
            "tinfo": w[o_desc + 4 * ntmax:o_desc + 8 * ntmax].view(ntmax, 4).tolist(),
            "wdesc": w[o_desc + 8 * ntmax:o_desc + 12 * ntmax].view(ntmax, 4).tolist(),
            "wids": w[o_desc + 12 * ntmax:o_desc + 44 * ntmax].view(ntmax, 32).tolist(),
            "cum": w[o_desc + 44 * ntmax:o_desc + 44 * ntmax + batch + 1].tolist()}


def paged_decode(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                 block_table: torch.Tensor, seq_lens: torch.Tensor, max_seq_len: int,
                 softmax_scale: float, out: torch.Tensor | None = None,
                 lse: torch.Tensor | None = None, kv_cache_dtype: str | int = "auto",
                 k_scale: torch.Tensor | None = None, v_scale: torch.Tensor | None = None,
                 num_splits: int = 0, variant: int = VARIANT_AUTO,
                 workspace: DecodeWorkspace | None = None, planned: bool = False,
                 early_kv_read: bool = False, early_q_read: bool = False) -> torch.Tensor:
    """Paged decode attention (query length 1). Writes `out` in place and returns it.

    planned: `decode_plan` was run for this step on `workspace` (skips the per-call plan launch).
    early_q_read: see B200ATTN_DECODE_EARLY_Q_READ (q complete before the preceding kernel started).
    early_kv_read: see B200ATTN_DECODE_EARLY_KV_READ in include/b200attn.h (the caller's cache write of this
    step only touched the last token of each sequence)."""
    lib = _lib.load()
    if out is None:
        out = torch.empty_like(q)
    batch, hq, d = q.shape
    if batch == 0:
        return out
    if workspace is None:
        if planned:
            raise ValueError("planned=True needs the workspace decode_plan() was run on")
        workspace = DecodeWorkspace(q.device)
    if num_splits != 1 or variant in (VARIANT_STREAMK, VARIANT_SK2, VARIANT_SK3):
        s = num_splits if num_splits > 0 else decode_num_splits(
            batch, hq, k_cache.shape[2], max_seq_len)
        workspace.ensure(decode_workspace_bytes(batch, hq, d, s))
    flags = ((DECODE_PLANNED if planned else 0) | (DECODE_EARLY_KV_READ if early_kv_read else 0) |
             (DECODE_EARLY_Q_READ if early_q_read else 0))
    args = _decode_args(q, k_cache, v_cache, out, lse, block_table, seq_lens, max_seq_len,
                        softmax_scale, kv_cache_dtype, k_scale, v_scale, num_splits, variant, workspace, flags)
    check("b200attn_paged_decode", lib.b200attn_paged_decode(*args))
    return out


def paged_decode_static(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                        block_table: torch.Tensor, seq_lens: torch.Tensor, max_seq_len: int,
                        softmax_scale: float, out: torch.Tensor, lse: torch.Tensor | None = None,
                        num_splits: int = 0) -> torch.Tensor:
    """`paged_decode` through b200attn_paged_decode_static: fixed-split schedule with the partial rows in the
    library's own 8 MiB device array — no workspace, and no allocation on this side either (`out` is the caller's).
    bf16 caches; batch * Hq * num_splits <= _lib.STATIC_MAX_ROWS. Calls that overlap on the GPU race on the array."""
    lib = _lib.load()
    if q.shape[0] == 0:
        return out
    a = _decode_args(q, k_cache, v_cache, out, lse, block_table, seq_lens, max_seq_len, softmax_scale,
                     "auto", None, None, num_splits, VARIANT_STREAM, None)
    # drop k_scale, v_scale, scale_numel, variant, flags, workspace, bytes
    check("b200attn_paged_decode_static", lib.b200attn_paged_decode_static(*a[:22], a[25], a[30]))
    return out


def paged_decode_partial(q, k_cache, v_cache, block_table, seq_lens, max_seq_len, softmax_scale,
                         out, num_splits: int, workspace: DecodeWorkspace,
                         lse=None, kv_cache_dtype="auto", k_scale=None, v_scale=None,
                         variant: int = VARIANT_AUTO, planned: bool = False, early_kv_read: bool = False,
                         early_q_read: bool = False) -> None:
    """First launch only (profiling sweeps): split-K partial kernel."""
    lib = _lib.load()
    workspace.ensure(decode_workspace_bytes(q.shape[0], q.shape[1], q.shape[2], num_splits))
    flags = ((DECODE_PLANNED if planned else 0) | (DECODE_EARLY_KV_READ if early_kv_read else 0) |
             (DECODE_EARLY_Q_READ if early_q_read else 0))
    args = _decode_args(q, k_cache, v_cache, out, lse, block_table, seq_lens, max_seq_len,
                        softmax_scale, kv_cache_dtype, k_scale, v_scale, num_splits, variant, workspace, flags)
    check("b200attn_paged_decode_partial", lib.b200attn_paged_decode_partial(*args))


def lse_reduce(out: torch.Tensor, seq_lens: torch.Tensor, max_seq_len: int, num_splits: int,
               workspace: DecodeWorkspace, lse: torch.Tensor | None = None, num_kv_heads: int | None = None,
               variant: int = VARIANT_AUTO) -> None:
    """Second launch only (profiling sweeps): merge the per-split partials by their LSE."""
    lib = _lib.load()
    _need_cuda(out, seq_lens, lse)
    batch, hq, d = out.shape
    check("b200attn_lse_reduce", lib.b200attn_lse_reduce(
        out.data_ptr(), out.stride(0), lse.data_ptr() if lse is not None else None,
        seq_lens.data_ptr(), batch, int(max_seq_len), hq, num_kv_heads if num_kv_heads is not None else hq, d,
        int(num_splits), int(variant), workspace.ptr, workspace.nbytes, _stream()))


class PrefillStaging:
    """Caller-owned bf16 staging of the fp8 pages a step's prefill sequences own (b200attn_paged_prefill_fp8_staged):
    one buffer serves every layer of a step in turn; grown on demand, outside graphs."""

    def __init__(self, device: torch.device | str = "cuda"):
        self.device = torch.device(device)
        self.buf: torch.Tensor | None = None

    def ensure(self, nbytes: int) -> None:
        if self.buf is None or self.buf.numel() < nbytes:
            if torch.cuda.is_current_stream_capturing():
                raise RuntimeError("fp8 prefill staging too small and cannot grow during graph capture")
            self.buf = None  # release before the larger allocation
            self.buf = torch.empty(int(nbytes * 1.25) + 1024, dtype=torch.uint8, device=self.device)

    @property
    def ptr(self) -> int:
        return self.buf.data_ptr()

    @property
    def nbytes(self) -> int:
        return self.buf.numel()


def prefill_staging_bytes(batch: int, max_seq_len: int, total_kv_tokens: int, num_kv_heads: int, head_size: int,
                          block_size: int) -> int:
    return int(_lib.load().b200attn_prefill_staging_bytes(int(batch), int(max_seq_len), int(total_kv_tokens),
                                                          int(num_kv_heads), int(head_size), int(block_size)))


def _fp8_prefill_staged(num_tokens: int, batch: int, group: int) -> bool:
    """Whether an fp8 prefill call takes the tcgen05 kernel through the bf16 staging.

    Staging moves 5 B per cached element of every sequence of the call; the tcgen05 kernel saves about
    (1/200 - 1/750) ps per flop against the mma.sync fp8 kernel (profiles/r2_prefill_fp8_staged.jsonl). With
    flops ~ 512 * group * new tokens * context / 2 per KV head the two meet at about 210 = group * new tokens per
    sequence: long prompts and chunked prefill gain 2-3.7x, short extends over long contexts (speculative decoding:
    a handful of new tokens per sequence) would lose — those stay on the mma.sync kernel.
    B200ATTN_PREFILL_FP8_STAGED: 0 = never, force = always, otherwise the rule above; B200ATTN_PREFILL_UMMA=0 (the
    switch that also sends bf16 caches to the mma.sync kernel) turns it off too."""
    import os
    mode = os.environ.get("B200ATTN_PREFILL_FP8_STAGED", "1").lower()
    if os.environ.get("B200ATTN_PREFILL_UMMA", "1")[:1] in ("0", "n", "N", "f", "F") or mode in ("0", "n", "no", "f", "false", "off"):
        return False
    return mode == "force" or num_tokens * group >= FP8_STAGED_MIN_WORK * batch


FP8_STAGED_MIN_WORK = 256  # group x new tokens per sequence from which the staged fp8 prefill pays (break-even ~210)


def paged_prefill(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                  block_table: torch.Tensor, seq_lens: torch.Tensor, query_start_loc: torch.Tensor,
                  max_query_len: int, softmax_scale: float, out: torch.Tensor | None = None,
                  kv_cache_dtype: str | int = "auto", k_scale=None, v_scale=None,
                  staging: PrefillStaging | None = None, max_seq_len: int = 0,
                  total_kv_tokens: int = 0) -> torch.Tensor:
    """Causal varlen attention over the paged cache (prefill / extend / mixed batches).

    staging + max_seq_len (host bound of seq_lens; total_kv_tokens = host bound of their sum, 0 if unknown): fp8
    caches are staged as bf16 pages and served by the tcgen05 kernel (b200attn_paged_prefill_fp8_staged) when the
    call has enough new tokens per sequence to pay for the staging (`_fp8_prefill_staged`); otherwise, and without a
    staging buffer, fp8 caches run on the mma.sync kernel."""
    lib = _lib.load()
    _need_cuda(q, k_cache, v_cache, block_table, seq_lens, query_start_loc)
    if out is None:
        out = torch.empty_like(q)
    code = kv_dtype_code(kv_cache_dtype)
    fp8 = code == KV_FP8_E4M3
    if q.dtype != torch.bfloat16 or q.ndim != 3 or q.stride(2) != 1 or q.stride(1) != q.shape[2]:
        raise ValueError("q must be bf16 [T, Hq, d] with contiguous heads")
    if fp8 and k_cache.dtype == torch.float8_e4m3fn:
        k_cache, v_cache = k_cache.view(torch.uint8), v_cache.view(torch.uint8)
    sb, ss, sh = _cache_strides(k_cache, v_cache)
    nb, bs, hkv, d = k_cache.shape
    if query_start_loc.dtype != torch.int32 or seq_lens.dtype != torch.int32:
        raise ValueError("query_start_loc/seq_lens must be int32")
    batch = query_start_loc.numel() - 1
    ks, vs, sn = _scales(k_scale, v_scale, fp8)
    if (fp8 and staging is not None and max_seq_len > 0 and batch > 0
            and _fp8_prefill_staged(q.shape[0], batch, q.shape[1] // hkv)):
        staging.ensure(prefill_staging_bytes(batch, max_seq_len, total_kv_tokens, hkv, d, bs))
        off = (-staging.ptr) % 1024  # the library wants the staging 1 KiB aligned
        check("b200attn_paged_prefill_fp8_staged", lib.b200attn_paged_prefill_fp8_staged(
            q.data_ptr(), q.stride(0), k_cache.data_ptr(), v_cache.data_ptr(), sb, ss, sh, nb,
            out.data_ptr(), out.stride(0), block_table.data_ptr(), block_table.stride(0),
            seq_lens.data_ptr(), query_start_loc.data_ptr(), batch, q.shape[0], int(max_query_len),
            int(max_seq_len), int(total_kv_tokens), q.shape[1], hkv, d, bs, float(softmax_scale), ks, vs, sn,
            staging.ptr + off, staging.nbytes - off, _stream()))
        return out
    check("b200attn_paged_prefill", lib.b200attn_paged_prefill(
        q.data_ptr(), q.stride(0), k_cache.data_ptr(), v_cache.data_ptr(), sb, ss, sh, nb,
        out.data_ptr(), out.stride(0), block_table.data_ptr(), block_table.stride(0),
        seq_lens.data_ptr(), query_start_loc.data_ptr(), batch, q.shape[0], int(max_query_len),
        q.shape[1], hkv, d, bs, float(softmax_scale), code, ks, vs, sn, _stream()))
    return out


def gather_kv(cache: torch.Tensor, block_ids: torch.Tensor, seq_len: int) -> torch.Tensor:
    """dst[s, h, :] = cache[block_ids[s // bs], s % bs, h, :] for s < seq_len (cache dtype kept)."""
    lib = _lib.load()
    _need_cuda(cache, block_ids)
    if cache.ndim != 4 or cache.stride(3) != 1:
        raise ValueError("cache must be a [nb, bs, Hkv, d] view with contiguous d")
    if block_ids.dtype != torch.int32 or not block_ids.is_contiguous():
        raise ValueError("block_ids must be contiguous int32")
    nb, bs, hkv, d = cache.shape
    code = KV_FP8_E4M3 if cache.element_size() == 1 else KV_BF16
    dst = torch.empty((seq_len, hkv, d), dtype=cache.dtype, device=cache.device)
    check("b200attn_gather_kv", lib.b200attn_gather_kv(
        cache.data_ptr(), cache.stride(0), cache.stride(1), cache.stride(2), nb,
        block_ids.data_ptr(), block_ids.numel(), int(seq_len), hkv, d, bs, code, dst.data_ptr(),
        _stream()))
    return dst
