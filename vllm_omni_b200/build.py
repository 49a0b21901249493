"""In-tree build of libb200attn.so (the C-ABI CUDA library) with nvcc for sm_100a.

The library is plain CUDA runtime code (no torch, no pybind): `nvcc -shared` of csrc/*.cu.
It is loaded with ctypes by `vllm_omni_b200._lib`. The built .so is git-ignored but travels to
the GPU box with the repo snapshot.
"""
from __future__ import annotations

import os
import shutil
import subprocess
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
CSRC = PKG_DIR / "csrc"
LIB_PATH = PKG_DIR / "libb200attn.so"
SOURCES = ["api.cu", "paged_decode.cu", "paged_decode_stream.cu", "paged_decode_streamk.cu", "paged_decode_sk2.cu", "paged_decode_sk3.cu", "cache_ops.cu", "paged_prefill.cu", "rope_cache.cu", "peer_allreduce.cu", "paged_prefill_umma.cu"]
HEADERS = ["ptx.cuh", "umma.cuh", "decode_common.cuh", "sk_math.cuh", "kernels.h", "../../include/b200attn.h"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "--expt-relaxed-constexpr",
    "-Xcompiler", "-fPIC,-fvisibility=hidden",
    "-Xptxas", "-v",
    "-shared",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found (set NVCC=/path/to/nvcc)")


def sources() -> list[Path]:
    return [CSRC / s for s in SOURCES if (CSRC / s).exists()]


def is_stale() -> bool:
    if not LIB_PATH.exists():
        return True
    t = LIB_PATH.stat().st_mtime
    deps = sources() + [CSRC / h for h in HEADERS]
    return any(d.exists() and d.stat().st_mtime > t for d in deps)


def build_library(force: bool = False, verbose: bool = False) -> Path:
    """Compile csrc/*.cu into libb200attn.so; returns its path. Raises on compiler errors."""
    if not force and not is_stale():
        return LIB_PATH
    cmd = [_nvcc(), *NVCC_FLAGS, "-I", str(PKG_DIR.parent / "include"), "-I", str(CSRC),
           "-o", str(LIB_PATH), *map(str, sources())]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    log = proc.stdout + proc.stderr
    (PKG_DIR / "build.log").write_text(" ".join(cmd) + "\n" + log)
    if proc.returncode != 0:
        raise RuntimeError(f"nvcc failed ({proc.returncode}):\n{log[-4000:]}")
    if verbose:
        print(log)
    return LIB_PATH


if __name__ == "__main__":
    print(build_library(force=True, verbose=True))
