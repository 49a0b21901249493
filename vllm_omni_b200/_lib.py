"""ctypes binding of libb200attn.so — the only caller of the C ABI in include/b200attn.h.

There is deliberately no fallback: if the library is missing or a symbol is absent the import
of the product path fails loudly (`B200AttnLibraryError`). Nothing here touches the oracle.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
LIB_PATH = PKG_DIR / "libb200attn.so"

ABI_VERSION = 10
PEER_FLAG_BYTES = 8192
AR_AUTO, AR_ONESHOT, AR_TWOSHOT, AR_MULTIMEM, AR_PUSH, AR_PUSH2 = 0, 1, 2, 3, 4, 5
KV_BF16 = 0
KV_FP8_E4M3 = 1
VARIANT_AUTO = 0
VARIANT_TMA = 1
VARIANT_CPASYNC = 2
VARIANT_STREAM = 3
VARIANT_STREAMK = 4
VARIANT_SK2 = 5
VARIANT_SK3 = 6
DECODE_PLANNED = 1
DECODE_EARLY_KV_READ = 2
DECODE_EARLY_Q_READ = 4
STATIC_SCRATCH_BYTES = 8388608   # b200attn_paged_decode_static: the library's own scratch array
STATIC_MAX_ROWS = 16000          # ... and the partial rows (batch x query heads x splits) it holds

ERRORS = {1: "EINVAL", 2: "EUNSUPPORTED", 3: "EWORKSPACE", 4: "EDRIVER"}


class B200AttnLibraryError(RuntimeError):
    """libb200attn.so is missing, stale or incomplete."""


class B200AttnError(RuntimeError):
    """A C-ABI call returned non-zero (negative: argument rejected, positive: cudaError_t)."""

    def __init__(self, fn: str, code: int, msg: str):
        self.fn, self.code = fn, code
        kind = ERRORS.get(-code, f"cudaError {code}") if code < 0 else f"cudaError {code}"
        super().__init__(f"{fn} failed [{kind}]: {msg}")


_vp, _i32, _i64, _f32, _sz = C.c_void_p, C.c_int32, C.c_int64, C.c_float, C.c_size_t

# name -> (restype, argtypes); mirrors include/b200attn.h one to one.
_CACHE_ARGS = [_i64, _i64, _i64]  # stride_block, stride_slot, stride_head
_DECODE_ARGS = [
    _vp, _i64,                       # q, q_stride_tok
    _vp, _vp, *_CACHE_ARGS, _i64,    # k_cache, v_cache, strides, num_blocks
    _vp, _i64,                       # out, out_stride_tok
    _vp,                             # lse
    _vp, _i64,                       # block_table, stride
    _vp, _i32, _i32,                 # seq_lens, batch, max_seq_len
    _i32, _i32, _i32, _i32,          # Hq, Hkv, head_size, block_size
    _f32, _i32,                      # softmax_scale, kv_dtype
    _vp, _vp, _i32,                  # k_scale, v_scale, scale_numel
    _i32, _i32, _i32,                # num_splits, variant, flags
    _vp, _sz, _vp,                   # workspace, bytes, stream
]
SIGNATURES: dict[str, tuple] = {
    "b200attn_abi_version": (C.c_int, []),
    "b200attn_last_error": (C.c_char_p, []),
    "b200attn_num_sms": (C.c_int, []),
    "b200attn_reshape_and_cache": (C.c_int, [
        _vp, _vp, _i64, _i64, _i64, _i64,
        _vp, _vp, *_CACHE_ARGS,
        _vp, _i32, _i32, _i32, _i32, _i64, _i32,
        _vp, _vp, _i32, _vp]),
    "b200attn_rope_and_cache": (C.c_int, [
        _vp, _i64, _i32,                 # query, token stride, Hq
        _vp, _vp, _i64, _i64, _i64, _i64,  # key, value, their token/head strides
        _vp, _i32, _i64,                 # positions, rows (1|3), row stride
        _vp, _i64, _i64,                 # cos_sin_cache, row stride, max_position
        _i32, _i32,                      # rot_dim, is_neox
        C.POINTER(_i32), _i32,           # mrope_section (host), interleaved
        _vp, _vp, *_CACHE_ARGS,          # caches + strides
        _vp, _i32, _i32,                 # slot_mapping, num_tokens, num_slot_tokens
        _i32, _i32, _i32, _i64, _i32,    # Hkv, d, block_size, num_blocks, kv_dtype
        _vp, _vp, _i32, _vp]),
    "b200attn_decode_num_splits": (_i32, [_i32, _i32, _i32, _i32, _i32]),
    "b200attn_decode_workspace_bytes": (_sz, [_i32, _i32, _i32, _i32]),
    "b200attn_workspace_init": (C.c_int, [_vp, _sz, _vp]),
    "b200attn_decode_plan": (C.c_int, [_vp, _i32, _i32, _i32, _vp, _i64, _i32, _vp, _sz, _vp]),
    "b200attn_debug_set_trace": (None, [_vp]),
    "b200attn_debug_set_team": (None, [C.c_int32]),
    "b200attn_debug_map_encodes": (C.c_uint64, []),
    "b200attn_debug_plan_offset": (C.c_size_t, [C.c_int32, C.c_int32]),
    "b200attn_debug_sk_schedule": (None, [_i64, _i32, _i32, C.POINTER(_i64)]),
    "b200attn_paged_decode": (C.c_int, _DECODE_ARGS),
    "b200attn_paged_decode_partial": (C.c_int, _DECODE_ARGS),
    "b200attn_static_scratch_bytes": (_sz, []),
    # _DECODE_ARGS without the scales, variant, flags and workspace
    "b200attn_paged_decode_static": (C.c_int, [*_DECODE_ARGS[:-9], _i32, _vp]),
    "b200attn_lse_reduce": (C.c_int, [
        _vp, _i64, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _vp, _sz, _vp]),
    "b200attn_paged_prefill": (C.c_int, [
        _vp, _i64,
        _vp, _vp, *_CACHE_ARGS, _i64,
        _vp, _i64,
        _vp, _i64,
        _vp, _vp,
        _i32, _i32, _i32,
        _i32, _i32, _i32, _i32,
        _f32, _i32,
        _vp, _vp, _i32, _vp]),
    "b200attn_prefill_staging_bytes": (_sz, [_i32, _i32, _i64, _i32, _i32, _i32]),
    "b200attn_paged_prefill_fp8_staged": (C.c_int, [
        _vp, _i64,
        _vp, _vp, *_CACHE_ARGS, _i64,
        _vp, _i64,
        _vp, _i64,
        _vp, _vp,
        _i32, _i32, _i32,
        _i32, _i64,                      # max_seq_len, total_kv_tokens
        _i32, _i32, _i32, _i32,
        _f32,
        _vp, _vp, _i32,
        _vp, _sz, _vp]),                 # staging, bytes, stream
    "b200attn_gather_kv": (C.c_int, [
        _vp, *_CACHE_ARGS, _i64, _vp, _i32, _i32, _i32, _i32, _i32, _i32, _vp, _vp]),
    "b200attn_peer_allreduce_algo": (_i32, [_i64, _i32, _i32, _i64]),
    "b200attn_peer_allreduce": (C.c_int, [
        C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp),   # peer_in, peer_scratch, peer_flags (host arrays)
        C.POINTER(_vp), _i64,                             # peer_push (host array or NULL), push_slot_bytes
        _vp, _vp, _vp,                                    # out, mc_in, mc_out
        _i64, _i32, _i32, _i32, _i32,                     # numel, world, rank, algo, max_blocks
        C.c_uint64, _vp]),                                # timeout_ns, stream
    "b200attn_peer_exchange": (C.c_int, [
        C.POINTER(_vp), C.POINTER(_vp), _vp,              # peer_rows, peer_flags (host arrays), local_rows
        _i32, _i32, _i32, _i32, _i32,                     # group_size, group_rank, first_row, num_rows, rows_per_rank
        _i64, _i32, C.c_uint64, _vp]),                    # row_bytes, scatter, timeout_ns, stream
    "b200attn_peer_alloc": (C.c_int, [_sz, C.POINTER(_vp), C.c_char_p]),
    "b200attn_peer_open": (C.c_int, [C.c_char_p, C.POINTER(_vp)]),
    "b200attn_peer_close": (C.c_int, [_vp]),
    "b200attn_peer_free": (C.c_int, [_vp]),
}

_lib: C.CDLL | None = None


def load(path: str | os.PathLike | None = None) -> C.CDLL:
    """Load (once) and type the library. Raises B200AttnLibraryError if it cannot be used."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = Path(path) if path is not None else LIB_PATH
    if not p.exists():
        raise B200AttnLibraryError(
            f"{p} not found — build it with `python -m vllm_omni_b200.build` "
            "(nvcc, sm_100a). There is no CPU or PyTorch fallback for this path.")
    try:
        lib = C.CDLL(str(p))
    except OSError as e:  # pragma: no cover - depends on the box
        raise B200AttnLibraryError(f"cannot load {p}: {e}") from e
    for name, (res, args) in SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as e:
            raise B200AttnLibraryError(f"{p} does not export {name}") from e
        fn.restype, fn.argtypes = res, args
    got = lib.b200attn_abi_version()
    if got != ABI_VERSION:
        raise B200AttnLibraryError(f"{p} has ABI {got}, this package needs {ABI_VERSION}; rebuild")
    if path is None:
        _lib = lib
    return lib


def check(fn: str, rc: int) -> None:
    if rc != 0:
        msg = load().b200attn_last_error()
        raise B200AttnError(fn, rc, msg.decode(errors="replace") if msg else "")
