"""Stream-K schedule (static chunks + claimed dynamic chunks): the exported host evaluation of the
function the kernels use must tile the stream exactly and stay inside the workspace bound."""
import ctypes as C

import pytest

from vllm_omni_b200 import _lib

W = 12            # kSkWarps
DYN_PER_WARP = 4  # kSkDynPerWarp


def sched(lib, total, nw, shift):
    out = (C.c_int64 * 5)()
    lib.b200attn_debug_sk_schedule(total, nw, shift, out)
    return dict(S=out[0], ts=out[1], cd=out[2], nd=out[3], dyn=bool(out[4]))


@pytest.mark.parametrize("shift", [0, 1, 2, 3, 4])
@pytest.mark.parametrize("nw", [12, 148 * W, 132 * W, 7])
def test_schedule_partitions_the_stream(lib, nw, shift):
    totals = [0, 1, 2, nw - 1, nw, nw + 1, 3 * nw + 5, 37 * nw - 3, 65536, 262144, 131072, 64, 1000003]
    for total in totals:
        s = sched(lib, total, nw, shift)
        assert 0 <= s["S"] <= total and s["ts"] >= 1 and s["cd"] >= 2
        # static chunks cover [0, S) with at most nw chunks of ts tiles
        assert s["S"] <= nw * s["ts"]
        # dynamic chunks cover [S, total) exactly
        assert (s["nd"] - 1) * s["cd"] < total - s["S"] <= s["nd"] * s["cd"] or (s["nd"] == 0 and total == s["S"])
        # workspace bound: at most DYN_PER_WARP dynamic chunks per warp (+1 for rounding)
        assert s["nd"] <= DYN_PER_WARP * nw + 1
        if shift == 0:
            assert not s["dyn"] and s["S"] == total
        if not s["dyn"]:
            assert s["S"] == total or s["nd"] == 0


def test_default_share_at_cfg3(lib):
    # cfg3: 64 seqs x 4 kv heads x 256 tiles on 148 x 12 warps: 37 tiles per warp, 4 of them dynamic
    s = sched(lib, 64 * 4 * 256, 148 * W, 3)
    assert s == dict(S=148 * W * 33, ts=33, cd=2, nd=(65536 - 148 * W * 33 + 1) // 2, dyn=True)
