"""GPU: the backend as a drop-in inside vLLM's own engine.

A random-init Qwen3-style decoder (16q/2kv heads, d=128, 4 layers; no checkpoint, no tokenizer) is
served by vLLM with `attention_backend=CUSTOM` — through `GPUModelRunner`, the metadata builder,
`Attention.forward`, `do_kv_cache_update` and `forward`, with chunked prefill of ragged prompts and
batched decode — and its greedy tokens / logprobs are compared with vLLM's stock backend on the
same weights and prompts. Each engine runs in its own subprocess (scripts/e2e_vllm.py)."""
import json
import subprocess
import sys
from pathlib import Path

import pytest

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parents[1]
STOCK = ["FLASH_ATTN", "FLASHINFER", "TRITON_ATTN"]


def run_engine(tmp_path, name, *args):
    out = tmp_path / f"{name}.json"
    p = subprocess.run([sys.executable, str(ROOT / "scripts" / "e2e_vllm.py"), *args, "--out", str(out)],
                       capture_output=True, text=True, timeout=900, cwd=str(ROOT))
    if p.returncode != 0 or not out.exists():
        return None, (p.stdout + p.stderr)[-3000:]
    return json.loads(out.read_text()), ""


def stock_reference(tmp_path, *args):
    for be in STOCK:
        ref, _ = run_engine(tmp_path, f"ref_{be}", "--backend", be, *args)
        if ref is not None:
            return ref
    pytest.skip("no stock vLLM attention backend runs this configuration on this box")


def logprob_floor(tmp_path, ref, *args):
    """How far two STOCK backends sit from each other on the same weights and prompts (max |logprob difference|
    over the tokens they agree on): the noise floor of a bf16 model under another summation order. None if only
    one stock backend runs here."""
    for be in STOCK:
        if be == ref["backend"]:
            continue
        other, _ = run_engine(tmp_path, f"floor_{be}", "--backend", be, *args)
        if other is not None:
            worst = 0.0
            for ra, rb in zip(ref["results"], other["results"]):
                for i, (ta, tb) in enumerate(zip(ra["tokens"], rb["tokens"])):
                    if ta != tb:
                        break
                    worst = max(worst, abs(ra["logprobs"][i] - rb["logprobs"][i]))
            return worst
    return None


def compare(a, b, tol, min_equal):
    n = same = 0
    worst = 0.0
    for ra, rb in zip(a["results"], b["results"]):
        for i, (ta, tb) in enumerate(zip(ra["tokens"], rb["tokens"])):
            n += 1
            if ta != tb:
                break  # later tokens belong to different sequences
            same += 1
            worst = max(worst, abs(ra["logprobs"][i] - rb["logprobs"][i]))
    lp = [x for r in a["results"] for x in r["logprobs"]]
    assert max(lp) - min(lp) > 0.5, "degenerate (uniform) logits: the comparison would be vacuous"
    assert same >= min_equal * n, f"only {same}/{n} greedy tokens agree"
    assert worst <= tol, f"logprob difference {worst:.3f} > {tol}"
    return same, n, worst


@pytest.fixture(scope="module")
def vllm_ok():
    pytest.importorskip("vllm")


@pytest.mark.timeout(2400)
def test_custom_backend_matches_stock_backend_eager_and_graphs(lib, vllm_ok, tmp_path):
    got, err = run_engine(tmp_path, "custom_eager", "--backend", "CUSTOM")
    assert got is not None, err
    assert got["impl"] == "B200AttentionImpl"
    ref = stock_reference(tmp_path)
    assert ref["impl"] != "B200AttentionImpl"
    # tolerance on the chosen tokens' logprobs: 0.08, or 2.5 x the distance between two stock backends when that is
    # larger. A random-init bf16 model amplifies summation-order noise through its layers: two stock backends sit
    # 0.03-0.06 apart here, this backend 0.06 from FLASH_ATTN, while at the kernel level its decode error against the
    # oracle is the same as (or below) FlashAttention-2's on these very shapes (profiles/r2_kernel_error_vs_fa2.txt:
    # excess 5.5e-4 both, per-row 1.7e-3 vs 2.5e-3). The kernel-level bound is the 2e-3 oracle test; this test is about
    # the plumbing (a wrong slot, head, scale or block id moves logprobs by whole units and flips tokens).
    floor = logprob_floor(tmp_path, ref)
    tol = max(0.08, 2.5 * floor) if floor is not None else 0.08
    print(f"stock-vs-stock logprob distance {floor}, tolerance {tol:.3f}")
    print("eager:", compare(got, ref, tol=tol, min_equal=0.9))
    # decode replayed from CUDA graphs (FULL_DECODE_ONLY): no host sync / allocation in forward()
    got_g, err = run_engine(tmp_path, "custom_graphs", "--backend", "CUSTOM", "--graphs")
    assert got_g is not None, err
    print("graphs:", compare(got_g, ref, tol=tol, min_equal=0.9))


@pytest.mark.timeout(2400)
def test_custom_backend_fp8_kv_cache(lib, vllm_ok, tmp_path):
    got, err = run_engine(tmp_path, "custom_fp8", "--backend", "CUSTOM", "--kv-cache-dtype", "fp8")
    assert got is not None, err
    assert got["impl"] == "B200AttentionImpl" and got["kv_cache_dtype"] == "fp8"
    # fp8 tolerance: e4m3 keeps 3 mantissa bits; against the bf16-KV run of the same backend the chosen
    # tokens' logprobs move by a few 1e-2 and most greedy tokens survive
    ref, err = run_engine(tmp_path, "custom_bf16", "--backend", "CUSTOM")
    assert ref is not None, err
    compare(got, ref, tol=0.4, min_equal=0.6)
    # and against a stock backend with an fp8 cache, when one supports it here. Not a tight bound:
    # the stock fp8 paths also quantise Q / use other accumulation orders, so they sit as far from each
    # other as fp8 sits from bf16 (measured 0.21 vs 0.24); the tight fp8 bound is the kernel-level
    # test against the oracle on the SAME quantised cache (tests/test_gpu_decode.py).
    for be in ("TRITON_ATTN", "FLASHINFER"):
        stock, _ = run_engine(tmp_path, f"ref_fp8_{be}", "--backend", be, "--kv-cache-dtype", "fp8")
        if stock is not None:
            same, n, worst = compare(got, stock, tol=0.4, min_equal=0.6)
            print(f"fp8 vs {be}: {same}/{n} tokens, max logprob diff {worst:.3f}")
            break
