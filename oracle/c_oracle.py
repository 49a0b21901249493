"""ctypes binding of oracle/liboracle.so (plain-C CPU oracle) — TEST INFRASTRUCTURE ONLY.

See paged_attn_oracle.c for what each function restates. `build()` compiles the library with gcc
(OpenMP when the compiler has it); nothing in `vllm_omni_b200/` imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import shutil
import subprocess
from pathlib import Path

import numpy as np
import torch

HERE = Path(__file__).resolve().parent
SRC = HERE / "paged_attn_oracle.c"
LIB = HERE / "liboracle.so"
_BASE = ["-O2", "-fPIC", "-shared", "-fno-fast-math", "-ffp-contract=off"]


def build(force: bool = False) -> Path:
    if LIB.exists() and not force and LIB.stat().st_mtime >= SRC.stat().st_mtime:
        return LIB
    cands = [c for c in ("/usr/bin/gcc", shutil.which("gcc"), shutil.which("cc"),
                         os.environ.get("CC")) if c and os.path.exists(c)]
    errs = []
    for omp in (["-fopenmp"], []):
        for cc in cands:
            p = subprocess.run([cc, *_BASE, *omp, "-o", str(LIB), str(SRC), "-lm"],
                               capture_output=True, text=True)
            if p.returncode == 0:
                return LIB
            errs.append(f"{cc} {' '.join(omp)}: {p.stderr.strip()[-300:]}")
    raise RuntimeError("cannot build liboracle.so:\n" + "\n".join(errs))


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        l = C.CDLL(str(build()))
        vp, i32, i64, f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_float
        l.oracle_e4m3_encode.argtypes = [vp, vp, i64]
        l.oracle_e4m3_decode.argtypes = [vp, vp, i64]
        l.oracle_reshape_and_cache.argtypes = [vp, vp, i64, i64, i64, i64, vp, vp, i64, i64, i64,
                                               vp, i32, i32, i32, i32, i32, vp, vp, i32]
        l.oracle_paged_decode.argtypes = [vp, i64, vp, vp, i64, i64, i64, vp, vp, vp, i64, vp,
                                          i32, i32, i32, i32, i32, f32, i32, vp, vp, i32]
        for f in (l.oracle_e4m3_encode, l.oracle_e4m3_decode, l.oracle_reshape_and_cache,
                  l.oracle_paged_decode):
            f.restype = None
        _lib = l
    return _lib


def e4m3_encode(x: np.ndarray) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float32)
    out = np.empty(x.shape, dtype=np.uint8)
    lib().oracle_e4m3_encode(x.ctypes.data, out.ctypes.data, x.size)
    return out


def e4m3_decode(c: np.ndarray) -> np.ndarray:
    c = np.ascontiguousarray(c, dtype=np.uint8)
    out = np.empty(c.shape, dtype=np.float32)
    lib().oracle_e4m3_decode(c.ctypes.data, out.ctypes.data, c.size)
    return out


def _cpu(t: torch.Tensor | None):
    assert t is None or t.device.type == "cpu"
    return t


def reshape_and_cache(key, value, k_cache, v_cache, slot_mapping, fp8=False, k_scale=None,
                      v_scale=None) -> None:
    """In place on CPU tensors; caches are [nb,bs,Hkv,d] views (any strides, contiguous d)."""
    for t in (key, value, k_cache, v_cache, slot_mapping):
        _cpu(t)
    assert key.dtype == torch.bfloat16 and slot_mapping.dtype == torch.int64
    nb, bs, hkv, d = k_cache.shape
    ks = k_scale.contiguous() if fp8 else None
    vs = v_scale.contiguous() if fp8 else None
    lib().oracle_reshape_and_cache(
        key.data_ptr(), value.data_ptr(), key.stride(0), key.stride(1), value.stride(0),
        value.stride(1), k_cache.data_ptr(), v_cache.data_ptr(), k_cache.stride(0),
        k_cache.stride(1), k_cache.stride(2), slot_mapping.data_ptr(), slot_mapping.numel(), hkv, d,
        bs, 1 if fp8 else 0, ks.data_ptr() if fp8 else None, vs.data_ptr() if fp8 else None,
        ks.numel() if fp8 else 1)


def paged_decode(q, k_cache, v_cache, block_table, seq_lens, softmax_scale, fp8=False,
                 k_scale=None, v_scale=None, num_threads: int | None = None):
    """Returns (out f32 [B,Hq,d], lse f32 [B,Hq]); double accumulation, OpenMP over (b, h)."""
    for t in (q, k_cache, v_cache, block_table, seq_lens):
        _cpu(t)
    assert q.dtype == torch.bfloat16 and q.stride(2) == 1 and q.stride(1) == q.shape[2]
    assert block_table.dtype == torch.int32 and seq_lens.dtype == torch.int32
    B, hq, d = q.shape
    nb, bs, hkv, _ = k_cache.shape
    out = torch.empty(B, hq, d, dtype=torch.float32)
    lse = torch.empty(B, hq, dtype=torch.float32)
    ks = k_scale.contiguous() if fp8 else None
    vs = v_scale.contiguous() if fp8 else None
    if num_threads is not None:
        os.environ["OMP_NUM_THREADS"] = str(num_threads)
    lib().oracle_paged_decode(
        q.data_ptr(), q.stride(0), k_cache.data_ptr(), v_cache.data_ptr(), k_cache.stride(0),
        k_cache.stride(1), k_cache.stride(2), out.data_ptr(), lse.data_ptr(),
        block_table.data_ptr(), block_table.stride(0), seq_lens.contiguous().data_ptr(), B, hq, hkv,
        d, bs, float(softmax_scale), 1 if fp8 else 0, ks.data_ptr() if fp8 else None,
        vs.data_ptr() if fp8 else None, ks.numel() if fp8 else 1)
    return out, lse


if __name__ == "__main__":
    print(build(force=True))
