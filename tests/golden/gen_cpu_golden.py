"""Generates the CPU golden fixtures under tests/golden/ (run in the BUILD container only).

  kv_reader_bf16.npz / kv_reader_fp8.npz
      The reference's own KV-cache reader — `OmniKVTransferManager._extract_kv_cache` and
      `normalize_layer_kv` imported from /root/reference
      (vllm_omni/distributed/omni_connectors/kv_transfer_manager.py:786-862,
       utils/kv_utils.py:378-450) — run on a cache that the oracle's cache write filled in the
      B200 backend's physical layout. Stores the token-major K/V, slots, block ids and what the
      reference reader extracted. Pins: cache-write addressing + layout contract.
      Shapes follow the reference's fixture tests/distributed/omni_connectors/test_kv_flow.py:98-141
      (10 blocks, 4 heads, block_ids [1,3,5]) with block_size/head_size raised to the backend's 16/128.

/root/reference does not exist on the GPU box: tests read only the .npz files.
"""
import sys
import types
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
REF = "/root/reference"
# import the reader modules without executing vllm_omni/__init__.py (it needs `aenum`)
for name in ["vllm_omni", "vllm_omni.distributed", "vllm_omni.distributed.omni_connectors",
             "vllm_omni.distributed.omni_connectors.utils"]:
    m = types.ModuleType(name)
    m.__path__ = [f"{REF}/{name.replace('.', '/')}"]
    sys.modules[name] = m

from vllm_omni.distributed.omni_connectors.kv_transfer_manager import OmniKVTransferManager  # noqa: E402

from oracle import oracle  # noqa: E402
from vllm_omni_b200 import layout  # noqa: E402
from vllm_omni_b200.workloads import default_scales  # noqa: E402

OUT = Path(__file__).resolve().parent


def gen(fp8: bool):
    g = torch.Generator().manual_seed(1234 + fp8)
    nb, bs, hkv, d = 10, 16, 4, 128
    block_ids, seq_len = [1, 3, 5], 40
    k = torch.randn(seq_len, hkv, d, generator=g).to(torch.bfloat16)
    v = torch.randn(seq_len, hkv, d, generator=g).to(torch.bfloat16)
    pos = torch.arange(seq_len)
    slots = torch.tensor(block_ids)[pos // bs] * bs + pos % bs
    ks, vs = default_scales(hkv) if fp8 else (None, None)
    layers = []
    for _ in range(2):
        cache = layout.alloc_kv_cache(nb, bs, hkv, d, "fp8_e4m3" if fp8 else "auto", "cpu")
        oracle.ref_reshape_and_cache(k, v, cache[0], cache[1], slots, fp8, ks, vs)
        layers.append(cache)
    mgr = object.__new__(OmniKVTransferManager)  # the reference's own tests build it the same way
    data = OmniKVTransferManager._extract_kv_cache(mgr, "req-0", block_ids, seq_len, layers, bs,
                                                   "uint8" if fp8 else "bfloat16")
    ek = data.layer_blocks["key_cache"][0]
    ev = data.layer_blocks["value_cache"][0]
    assert ek.shape == (seq_len, hkv, d)
    as_np = (lambda t: t.view(torch.int16).numpy()) if not fp8 else (lambda t: t.numpy())
    np.savez_compressed(
        OUT / f"kv_reader_{'fp8' if fp8 else 'bf16'}.npz",
        key=k.view(torch.int16).numpy(), value=v.view(torch.int16).numpy(), slots=slots.numpy(),
        block_ids=np.array(block_ids, dtype=np.int32), seq_len=seq_len,
        shape=np.array([nb, bs, hkv, d]),
        k_scale=ks.numpy() if fp8 else np.zeros(0, np.float32),
        v_scale=vs.numpy() if fp8 else np.zeros(0, np.float32),
        extracted_key=as_np(ek), extracted_value=as_np(ev))
    print("wrote", "fp8" if fp8 else "bf16", ek.dtype, tuple(ek.shape))


if __name__ == "__main__":
    gen(False)
    gen(True)
