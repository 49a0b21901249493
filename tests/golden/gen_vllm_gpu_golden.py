#!/usr/bin/env python
"""Golden vectors for the ATTENTION OUTPUT of the path, produced by the reference's own stock code:
the kernels vLLM's FLASH_ATTN and FLASHINFER backends run for decode (what vllm-omni's AR stages use
before this repo's backend is registered) and vLLM's `reshape_and_cache_flash` for the slot writes.

Runs on a GPU box with vLLM installed (no /root/reference needed; the arithmetic of this path lives in
vLLM):

    python tests/golden/gen_vllm_gpu_golden.py gpurun_out/golden     # then copy *.npz to tests/golden/

Each file holds the complete inputs (q, paged cache as written by vLLM's own cache-write kernel, block
table, seq_lens, scale) and the outputs of every stock kernel that loaded, so the CPU suite can pin the
oracle (`oracle.ref_paged_decode`, the C oracle) against them without a GPU:
  out_fa2                   vllm.vllm_flash_attn.flash_attn_varlen_func (flash_attn.py:758-819 call shape)
  out_trtllm                flashinfer trtllm_batch_decode_with_kv_cache (flashinfer.py:1694-1822)
  cache_vllm                the cache after vLLM's reshape_and_cache_flash (vllm/_custom_ops.py:2591-2610)
bf16 arrays are stored as their int16 bit patterns.
"""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
import numpy as np
import torch

from tests.helpers import make_case

CASES = {
    # name: (seq_lens, hq, hkv, seed)  — thinker / talker head shapes of SURVEY §8, ragged lengths incl. 1 and
    # non-multiples of the block size
    "decode_32q4kv": ([130, 17, 1, 64], 32, 4, 11),
    "decode_28q4kv": ([96, 33], 28, 4, 12),
    "decode_16q2kv": ([48, 1, 15, 16, 150], 16, 2, 13),
}


def bits(t: torch.Tensor) -> np.ndarray:
    return t.detach().cpu().contiguous().view(torch.int16).numpy()


PREFILL_CASES = {
    # name: (seq_lens incl. the new tokens, new tokens per sequence, hq, hkv, seed): prefill, extend, single-token row
    "prefill_32q4kv": ([130, 17, 64], [40, 17, 1], 32, 4, 21),
    "prefill_16q2kv": ([48, 150], [48, 33], 16, 2, 22),
}


def prefill(out_dir: Path):
    """vLLM's FLASH_ATTN backend call for max_query_len > 1 (flash_attn.py:758-819) on the same kind of cache."""
    from vllm import _custom_ops as vops
    from vllm.vllm_flash_attn import flash_attn_varlen_func
    for name, (lens, q_lens, hq, hkv, seed) in PREFILL_CASES.items():
        c = make_case(lens, hq=hq, hkv=hkv, seed=seed)
        d, bs, nb = 128, c.block_size, c.num_blocks
        kv = torch.zeros(2, nb, bs, hkv, d, dtype=torch.bfloat16, device="cuda")
        one = torch.ones(1, dtype=torch.float32, device="cuda")
        vops.reshape_and_cache_flash(c.hist_k.cuda(), c.hist_v.cuda(), kv[0], kv[1], c.slots.cuda(), "auto", one, one)
        g = torch.Generator().manual_seed(seed + 100)
        q = torch.randn(sum(q_lens), hq, d, generator=g).to(torch.bfloat16)
        qsl = torch.tensor([0] + list(torch.tensor(q_lens).cumsum(0)), dtype=torch.int32)
        out = torch.empty_like(q, device="cuda")
        flash_attn_varlen_func(q.cuda(), kv[0], kv[1], max_seqlen_q=max(q_lens), cu_seqlens_q=qsl.cuda(),
                               max_seqlen_k=c.max_seq_len, seqused_k=c.seq_lens.cuda(), softmax_scale=c.scale,
                               causal=True, block_table=c.block_table.cuda(), out=out, fa_version=2)
        torch.cuda.synchronize()
        np.savez_compressed(out_dir / f"vllm_{name}.npz", q=bits(q), cache_vllm=bits(kv), block_table=c.block_table.numpy(),
                            seq_lens=c.seq_lens.numpy(), query_start_loc=qsl.numpy(), scale=np.float32(c.scale), hq=hq,
                            hkv=hkv, seed=seed, out_fa2=bits(out))
        print(f"{name}: wrote out_fa2")


FP8_CASES = {
    # name: (seq_lens, hq, hkv, seed, k_scale, v_scale) — scalar scales: what the stock fp8 decode can express
    "decode_fp8_32q4kv": ([130, 17, 1, 64, 300], 32, 4, 31, 0.5, 0.25),
    "decode_fp8_16q2kv": ([48, 1, 15, 16, 150], 16, 2, 32, 1.0, 2.0),
}


def fp8(out_dir: Path):
    """fp8-e4m3 KV cache with per-tensor scales: vLLM's fp8 `reshape_and_cache_flash` writes the cache bytes, FlashInfer's
    TRT-LLM-gen decode (what vLLM's FLASHINFER backend runs for an fp8 cache, flashinfer.py:1694-1822) reads them.
    Stored: the cache bytes (uint8) and the decode output, with the query dtype the stock kernel was given."""
    from vllm import _custom_ops as vops
    import flashinfer
    for name, (lens, hq, hkv, seed, ks, vs) in FP8_CASES.items():
        c = make_case(lens, hq=hq, hkv=hkv, seed=seed)
        d, bs, nb = 128, c.block_size, c.num_blocks
        kv = torch.zeros(2, nb, bs, hkv, d, dtype=torch.uint8, device="cuda")
        kst = torch.full((1,), ks, dtype=torch.float32, device="cuda")
        vst = torch.full((1,), vs, dtype=torch.float32, device="cuda")
        vops.reshape_and_cache_flash(c.hist_k.cuda(), c.hist_v.cuda(), kv[0], kv[1], c.slots.cuda(), "fp8", kst, vst)
        torch.cuda.synchronize()
        rec = dict(q=bits(c.q), cache_vllm_fp8=kv.cpu().numpy(), block_table=c.block_table.numpy(),
                   seq_lens=c.seq_lens.numpy(), scale=np.float32(c.scale), hq=hq, hkv=hkv, seed=seed,
                   k_scale=np.float32(ks), v_scale=np.float32(vs))
        q, bt, sl = c.q.cuda(), c.block_table.cuda(), c.seq_lens.cuda()
        ws = torch.zeros(128 * 1024 * 1024, dtype=torch.uint8, device="cuda")
        k = kv[0].permute(0, 2, 1, 3).contiguous().view(torch.float8_e4m3fn)   # HND
        v = kv[1].permute(0, 2, 1, 3).contiguous().view(torch.float8_e4m3fn)
        for qname, qq, qs in (("bf16", q, 1.0), ("fp8", q.to(torch.float8_e4m3fn), 1.0)):
            try:
                out = torch.empty_like(q)
                flashinfer.decode.trtllm_batch_decode_with_kv_cache(qq, (k, v), ws, bt, sl, c.max_seq_len,
                                                                    bmm1_scale=c.scale * ks * qs, bmm2_scale=vs, out=out,
                                                                    kv_layout="HND")
                torch.cuda.synchronize()
                rec[f"out_trtllm_q{qname}"] = bits(out)
            except Exception as e:  # noqa: BLE001
                print(f"{name}: trtllm-gen with {qname} q unavailable: {type(e).__name__}: {str(e)[:160]}")
        np.savez_compressed(out_dir / f"vllm_{name}.npz", **rec)
        print(f"{name}: wrote {sorted(k for k in rec if k.startswith('out_'))}")


def main(out_dir: Path):
    out_dir.mkdir(parents=True, exist_ok=True)
    assert torch.cuda.is_available()
    if "--prefill-only" in sys.argv:
        return prefill(out_dir)
    if "--fp8-only" in sys.argv:
        return fp8(out_dir)
    prefill(out_dir)
    fp8(out_dir)
    from vllm import _custom_ops as vops
    for name, (lens, hq, hkv, seed) in CASES.items():
        c = make_case(lens, hq=hq, hkv=hkv, seed=seed)
        d, bs, nb = 128, c.block_size, c.num_blocks
        # vLLM's own cache write into an NHD cache [2, nb, bs, Hkv, d]
        kv = torch.zeros(2, nb, bs, hkv, d, dtype=torch.bfloat16, device="cuda")
        one = torch.ones(1, dtype=torch.float32, device="cuda")
        vops.reshape_and_cache_flash(c.hist_k.cuda(), c.hist_v.cuda(), kv[0], kv[1], c.slots.cuda(), "auto", one, one)
        torch.cuda.synchronize()
        rec = dict(q=bits(c.q), cache_vllm=bits(kv), block_table=c.block_table.numpy(), seq_lens=c.seq_lens.numpy(),
                   scale=np.float32(c.scale), hq=hq, hkv=hkv, seed=seed)
        q, bt, sl = c.q.cuda(), c.block_table.cuda(), c.seq_lens.cuda()
        B = q.shape[0]
        cu_q = torch.arange(B + 1, dtype=torch.int32, device="cuda")
        try:
            from vllm.vllm_flash_attn import flash_attn_varlen_func
            for ver in (2,):   # v3 is the Hopper build (fatal 'no kernel image' on sm_100), v4 is not shipped for decode
                try:
                    out = torch.empty_like(q)
                    flash_attn_varlen_func(q, kv[0], kv[1], max_seqlen_q=1, cu_seqlens_q=cu_q,
                                           max_seqlen_k=c.max_seq_len, seqused_k=sl, softmax_scale=c.scale,
                                           causal=True, block_table=bt, out=out, fa_version=ver)
                    torch.cuda.synchronize()
                    rec[f"out_fa{ver}"] = bits(out)
                except Exception as e:  # noqa: BLE001
                    print(f"{name}: flash-attn v{ver} unavailable: {type(e).__name__}: {str(e)[:120]}")
        except Exception as e:  # noqa: BLE001
            print(f"{name}: vllm_flash_attn unavailable: {e}")
        try:
            import flashinfer
            ws = torch.zeros(128 * 1024 * 1024, dtype=torch.uint8, device="cuda")
            k = kv[0].permute(0, 2, 1, 3).contiguous()   # HND
            v = kv[1].permute(0, 2, 1, 3).contiguous()
            out = torch.empty_like(q)
            flashinfer.decode.trtllm_batch_decode_with_kv_cache(q, (k, v), ws, bt, sl, c.max_seq_len,
                                                                bmm1_scale=c.scale, bmm2_scale=1.0, out=out,
                                                                kv_layout="HND")
            torch.cuda.synchronize()
            rec["out_trtllm"] = bits(out)
        except Exception as e:  # noqa: BLE001
            print(f"{name}: flashinfer trtllm-gen unavailable: {type(e).__name__}: {str(e)[:120]}")
        np.savez_compressed(out_dir / f"vllm_{name}.npz", **rec)
        print(f"{name}: wrote {sorted(k for k in rec if k.startswith('out_'))}")


if __name__ == "__main__":
    main(Path(sys.argv[1] if len(sys.argv) > 1 and not sys.argv[1].startswith("--") else "gpurun_out/golden"))
