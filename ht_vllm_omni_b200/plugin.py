"""Plugin entry point: makes the B200 paged-attention backend selectable as `attention_backend: CUSTOM`.

Loaded through the entry-point groups `vllm_omni.general_plugins` (vllm_omni/plugins/__init__.py:14-18,61-74 — run in the
API process by prepare_engine_environment, engine/stage_init_utils.py:199-203, and in every worker by
OmniWorkerMixin.__init__, worker/mixins.py:6-13) and `vllm.general_plugins` (so EngineCore processes see it too).
Plugins "can be loaded multiple times ... must be idempotent" (plugins/__init__.py:62-65): register() only writes one
dict entry in vLLM's backend registry ($VLLM/v1/attention/backends/registry.py:203-255) and imports nothing heavy.
"""
from __future__ import annotations

BACKEND_PATH = "ht_vllm_omni_b200.backend.B200PagedAttentionBackend"


def register(override: str | None = None) -> str:
    """Register the backend under AttentionBackendEnum.CUSTOM (default) or in place of a built-in one
    (`override="FLASHINFER"` makes vLLM's automatic sm_100 choice, $VLLM/platforms/cuda.py:132-139, land on us)."""
    from vllm.v1.attention.backends.registry import AttentionBackendEnum, register_backend

    slot = AttentionBackendEnum[override] if override else AttentionBackendEnum.CUSTOM
    register_backend(slot, BACKEND_PATH)
    return slot.name


def register_omni_platform() -> None:
    """`vllm_omni.platform_plugins` hook: nothing to swap.  CudaOmniPlatform picks worker classes, stage configs and the
    diffusion attention backend only (vllm_omni/platforms/cuda/platform.py:24-81); the AR attention backend is chosen by
    vLLM's own selector ($VLLM/v1/attention/selector.py:111-117), which register() feeds.  Returns None = "no platform"."""
    return None
