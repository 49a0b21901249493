"""GPU: the backend inside vllm-omni itself. The UNMODIFIED reference package
(baseline/_ref, installed by scripts/install_reference.sh; skipped when it is absent) drives the run:

  * `OmniEngineArgs.__post_init__` calls the reference's plugin loader (vllm_omni/plugins/__init__.py:61-74), which
    finds this repo's `vllm_omni.general_plugins` entry point (pyproject.toml; in-tree dist-info for an un-installed
    checkout) and so registers `attention_backend: CUSTOM`;
  * the stage's engine_args get the backend the way the reference forces one for ROCm
    (vllm_omni/engine/stage_init_utils.py:282-293): `plugin.force_for_stage`;
  * the worker is `GPUARWorker`, whose `GPUARModelRunner.execute_model` (gpu_ar_model_runner.py:263-624) builds the
    attention metadata, and the reference's builder patch walks `attn_group.metadata_builders`
    (worker/gpu_model_runner.py:51-75) over `B200AttentionMetadataBuilder`;
  * the model is the reference's own Qwen2 decoder (qwen2_old.py: qkv -> rope -> self.attn -> o_proj).

Greedy tokens and logprobs must match the same run with a stock vLLM backend. Each engine runs in its own
subprocess (scripts/e2e_vllm.py --omni)."""
import pytest

from tests import ref_import
from tests.test_gpu_e2e_vllm import STOCK, compare, run_engine

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def omni_ok():
    pytest.importorskip("vllm")
    if not ref_import.available():
        pytest.skip("baseline/_ref (the reference package) is not installed in this checkout")


def stock_omni(tmp_path, *args):
    for be in STOCK:
        ref, _ = run_engine(tmp_path, f"omni_ref_{be}", "--omni", "--backend", be, *args)
        if ref is not None:
            return ref
    pytest.skip("no stock vLLM attention backend runs under the reference's runner on this box")


@pytest.mark.timeout(2400)
@pytest.mark.parametrize("graphs", [False, True])
def test_backend_under_the_reference_worker_and_model_runner(lib, omni_ok, tmp_path, graphs):
    extra = ("--graphs",) if graphs else ()
    got, err = run_engine(tmp_path, "omni_custom", "--omni", "--backend", "CUSTOM", *extra)
    assert got is not None, err
    assert got["impl"] == "B200AttentionImpl"
    assert got["worker"] == "GPUARWorker" and got["model_runner"] == "GPUARModelRunner"
    assert got["model_config"] == "OmniModelConfig" and got["omni_plugins_loaded"] is True
    assert got["metadata_builders"] == ["B200AttentionMetadataBuilder"]
    ref = stock_omni(tmp_path)
    assert ref["impl"] != "B200AttentionImpl" and ref["model_runner"] == "GPUARModelRunner"
    same, n, worst = compare(got, ref, tol=0.08, min_equal=0.9)
    print(f"vllm-omni runner, graphs={graphs}: {same}/{n} greedy tokens equal, max logprob difference {worst:.4f}; "
          f"stand-ins for missing imports: {got['stubbed_imports']}; vLLM version-skew adapters: "
          f"{got['vllm_version_skew_adapters']}")


@pytest.mark.timeout(2400)
def test_per_head_fp8_scales_through_the_reference_quant_routing(lib, omni_ok, tmp_path):
    """a checkpoint-style compressed-tensors KV-cache scheme (fp8, strategy attn_head) is routed by
    the reference's `ComponentQuantizationConfig` (vllm_omni/quantization/component_config.py:63-82: longest layer-name
    prefix; "model." -> the checkpoint's config, everything else -> none) to vLLM's KV-cache method, which gives every
    attention layer `[Hkv]` k/v scales; `B200AttentionImpl` must receive them as such (`scale_numel == Hkv`) and use
    the right head's scale: the two heads' scales differ by 16x (k 0.25 / 4.0, v 4.0 / 0.25), so indexing them wrongly
    destroys the output, while indexing them rightly leaves the usual fp8-vs-bf16 distance."""
    got, err = run_engine(tmp_path, "omni_perhead", "--omni", "--backend", "CUSTOM", "--kv-cache-dtype", "fp8",
                          "--kv-scales", "per-head")
    assert got is not None, err
    assert got["impl"] == "B200AttentionImpl" and got["model_runner"] == "GPUARModelRunner"
    assert got["quant_routing"] == {"attn": "CompressedTensorsConfig", "lm_head": "NoneType"}
    assert got["impl_scale_numel"] == 2 and got["kv_scales"] == {"k": [0.25, 4.0], "v": [4.0, 0.25]}
    ref, err = run_engine(tmp_path, "omni_bf16", "--omni", "--backend", "CUSTOM")
    assert ref is not None, err
    same, n, worst = compare(got, ref, tol=0.4, min_equal=0.6)
    print(f"per-head fp8 cache vs bf16 cache under the reference's runner: {same}/{n} greedy tokens, "
          f"max logprob difference {worst:.3f}")
