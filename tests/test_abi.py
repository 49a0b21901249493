"""CPU: the C-ABI library builds, loads without a GPU and exports every symbol the header declares."""
import ctypes
import re
from pathlib import Path

import pytest

from vllm_omni_b200 import _lib, build

ROOT = Path(__file__).resolve().parent.parent
HEADER = ROOT / "include" / "b200attn.h"


def declared_symbols() -> list[str]:
    text = re.sub(r"/\*.*?\*/", "", HEADER.read_text(), flags=re.S)
    return sorted(set(re.findall(r"\b(b200attn_\w+)\s*\(", text)))


def test_header_declares_the_expected_entry_points():
    syms = declared_symbols()
    for name in ("b200attn_reshape_and_cache", "b200attn_paged_decode", "b200attn_paged_decode_partial",
                 "b200attn_lse_reduce", "b200attn_decode_workspace_bytes", "b200attn_abi_version", "b200attn_decode_plan",
                 "b200attn_last_error", "b200attn_paged_prefill", "b200attn_gather_kv"):
        assert name in syms


def test_library_builds_and_exports_every_declared_symbol(lib):
    raw = ctypes.CDLL(str(build.LIB_PATH))
    for name in declared_symbols():
        assert hasattr(raw, name), f"{name} declared in b200attn.h but not exported"
    assert set(_lib.SIGNATURES) == set(declared_symbols())
    assert lib.b200attn_abi_version() == _lib.ABI_VERSION
    m = re.search(r"#define B200ATTN_ABI_VERSION (\d+)", HEADER.read_text())
    assert int(m.group(1)) == _lib.ABI_VERSION


def test_sass_is_blackwell_native():
    """The built library carries sm_100a code with TMA (UTMALDG), mbarrier (SYNCS), ldmatrix (LDSM),
    tensor-core MMA (HMMA), cp.async (LDGSTS) and tcgen05 (UTCHMMA = tcgen05.mma, UTCBAR = tcgen05.commit,
    LDTM / STTM = tcgen05.ld / st) instructions."""
    import shutil
    import subprocess
    if shutil.which("cuobjdump") is None:
        pytest.skip("cuobjdump not on PATH")
    build.build_library()
    out = subprocess.run(["cuobjdump", "-sass", str(build.LIB_PATH)], capture_output=True, text=True).stdout
    assert "sm_100a" in out
    for op in ("UTMALDG", "SYNCS", "LDSM", "HMMA", "LDGSTS", "MOVM", "UTCHMMA", "UTCBAR", "LDTM", "STTM"):
        assert op in out, f"{op} missing from SASS"


def test_split_heuristic_is_pure_host_arithmetic(lib):
    f = lib.b200attn_decode_num_splits
    # cfg3: 64 x 4 units of 256 tiles on 148 SMs must be split (256 units leave 40 SMs idle in wave 2)
    s = f(64, 32, 4, 4096, 148)
    assert s > 1
    items = 64 * 4 * s
    rounds = -(-items // 148)
    assert items / (rounds * 148) > 0.95
    # tiny work never splits below 128 tokens per split
    assert f(1, 16, 2, 100, 148) == 1
    assert f(1, 16, 2, 512, 148) <= 4
    # long context, small batch: many splits
    assert f(16, 32, 4, 32768, 148) >= 8
    # degenerate inputs
    assert f(0, 32, 4, 4096, 148) == 1 and f(64, 32, 4, 0, 148) == 1


def test_workspace_bytes_formula(lib):
    f = lib.b200attn_decode_workspace_bytes
    assert f(0, 32, 128, 4) == 0
    # max over the schedules; without a GPU the stream-K parts assume 148 SMs x 12 warps.
    # Every layout starts with a 64 KiB control area; then partial O rows and partial LSEs.
    ctrl = 65536
    rows = 64 * 32 * 8
    fixed = ctrl + rows * 128 * 4 + rows * 4 + 256
    # stream-K v1: two partial slots per chunk (one chunk per warp, +1), then the tiles-prefix plan
    sk_rows = 2 * (148 * 12 + 1) * 8

    def sk1(batch):
        return ctrl + sk_rows * 129 * 4 + ((batch + 1 + 3) // 4 * 4) * 4 + 256

    # stream-K v2 (default): warp partials, then the plan (header 16 | chunk starts warps + 1 | desc 4 x warps |
    # ids 32 x warps | cum batch + 1 | first/last chunk per item, items <= batch x query heads)
    def sk2(batch, hq):
        nw = 148 * 12
        plan = 16 + ((nw + 1 + 3) // 4 * 4) + 36 * nw + ((batch + 1 + 3) // 4 * 4) + 2 * batch * hq
        return ctrl + sk_rows * 129 * 4 + (plan + 4) * 4 + 256

    # cooperative stream-K (default), behind the v2 region: partial rows (2 parities x 2 per team, teams <= warps)
    # with their LSEs, then its plan (header 16 | team starts | team desc, team info, warp desc 4 x warps each | first block ids 32 x warps | cum)
    def sk3(batch, hq):
        nw = 148 * 12
        off = (sk2(batch, hq) + 255) // 256 * 256
        plan = 16 + ((nw + 1 + 3) // 4 * 4) + 44 * nw + ((batch + 1 + 3) // 4 * 4)
        return off + 2 * 2 * nw * 8 * 129 * 4 + (plan + 4) * 4 + 256

    assert f(64, 32, 128, 8) == max(fixed, sk1(64), sk3(64, 32))
    assert f(1, 8, 128, 1) == sk3(1, 8)


def test_zero_token_calls_are_noops_without_a_gpu(lib):
    # num_tokens == 0 / batch == 0 return before touching the device
    assert lib.b200attn_reshape_and_cache(None, None, 0, 0, 0, 0, None, None, 0, 0, 0, None, 0, 4,
                                          128, 16, 1, 0, None, None, 1, None) == 0
    rc = lib.b200attn_reshape_and_cache(None, None, 0, 0, 0, 0, None, None, 0, 0, 0, None, -1, 4,
                                        128, 16, 1, 0, None, None, 1, None)
    assert rc == -1 and b"num_tokens" in lib.b200attn_last_error()


def test_missing_library_fails_loudly(tmp_path):
    with pytest.raises(_lib.B200AttnLibraryError):
        _lib.load(tmp_path / "libb200attn.so")


def test_cpu_tensors_are_rejected_not_emulated(lib):
    import torch
    from vllm_omni_b200 import layout, ops
    cache = layout.alloc_kv_cache(4, 16, 2, 128, "auto", "cpu")
    k = torch.zeros(1, 2, 128, dtype=torch.bfloat16)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.reshape_and_cache(k, k, cache[0], cache[1], torch.zeros(1, dtype=torch.int64))
