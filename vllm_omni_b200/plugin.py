"""Registration hooks: make the B200 backend selectable as `attention_backend: CUSTOM`.

`register()` is the entry-point callable for the plugin groups that vllm-omni and vLLM load in
EVERY process (front-end, EngineCore, each TP worker):
  * `vllm_omni.general_plugins` — vllm_omni/plugins/__init__.py:14-18,61-74, executed from
    vllm_omni/worker/mixins.py:6-13 and vllm_omni/engine/arg_utils.py:154-156;
  * `vllm.general_plugins`      — vLLM's own loader.
It is idempotent (the loaders may call it several times per process) and honours VLLM_PLUGINS
through the loaders' allow-list. `force_for_stage()` mirrors the reference's ROCm precedent
(vllm_omni/engine/stage_init_utils.py:282-293) for callers that build engine_args themselves.
"""
from __future__ import annotations

BACKEND_PATH = "vllm_omni_b200.backend.B200AttentionBackend"


def register() -> None:
    from vllm.v1.attention.backends.registry import AttentionBackendEnum, register_backend

    register_backend(AttentionBackendEnum.CUSTOM, BACKEND_PATH)


def force_for_stage(engine_args: dict) -> dict:
    """Set `attention_backend: CUSTOM` on a stage's engine_args unless the user chose one."""
    register()
    if engine_args.get("attention_backend") is None:
        engine_args["attention_backend"] = "CUSTOM"
    return engine_args
