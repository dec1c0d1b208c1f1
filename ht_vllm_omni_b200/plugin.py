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
    (`override="FLASHINFER"` makes vLLM's automatic sm_100 choice, $VLLM/platforms/cuda.py:132-139, land on us), and -- when
    the reference is importable -- let its quantisation routing hand per-head KV scales to layers this backend serves."""
    from vllm.v1.attention.backends.registry import AttentionBackendEnum, register_backend

    slot = AttentionBackendEnum[override] if override else AttentionBackendEnum.CUSTOM
    register_backend(slot, BACKEND_PATH)
    install_per_head_kv_scales()
    return slot.name


def _per_head_method_for(layer, delegated):
    """The KV-cache quant method for `layer`, or None to keep `delegated` (what the component's vLLM config returned).
    Only FP8-KV Attention layers whose selected backend is ours are touched, and only where vLLM itself would have put a
    BaseKVCacheMethod (so checkpoints without KV scales, other layer types and other backends behave exactly as before)."""
    from .kv_scales import PerHeadKVCacheMethod

    try:
        from vllm.model_executor.layers.quantization.kv_cache import BaseKVCacheMethod
    except Exception:  # noqa: BLE001
        return None
    if not isinstance(delegated, BaseKVCacheMethod) or isinstance(delegated, PerHeadKVCacheMethod):
        return None
    if getattr(layer, "kv_cache_dtype", "auto") not in ("fp8", "fp8_e4m3"):
        return None
    backend = getattr(layer, "attn_backend", None)
    if backend is None or f"{backend.__module__}.{backend.__qualname__}" != BACKEND_PATH:
        return None
    tp, rank = 1, 0
    try:
        from vllm.distributed import get_tensor_model_parallel_rank, get_tensor_model_parallel_world_size

        tp, rank = get_tensor_model_parallel_world_size(), get_tensor_model_parallel_rank()
    except Exception:  # noqa: BLE001  (no TP group: a single-process engine)
        pass
    local = int(layer.num_kv_heads)
    # vLLM's head partitioning ($VLLM/model_executor/models/qwen3.py:78-95): Hkv / tp heads per rank, or ONE replicated head when
    # tp > Hkv -- in which case the model's head count is not recoverable from the layer and must come from its config
    total = getattr(layer, "total_num_kv_heads", None) or (local * tp if local > 1 or tp == 1 else None)
    if total is None:
        cfg = getattr(getattr(layer, "vllm_config", None), "model_config", None)
        total = cfg.get_total_num_kv_heads() if cfg is not None else local * tp
    return PerHeadKVCacheMethod(local, total_num_kv_heads=int(total), tp_size=tp, tp_rank=rank)


def install_per_head_kv_scales() -> bool:
    """Wraps ComponentQuantizationConfig.get_quant_method (vllm_omni/quantization/component_config.py:53-57) so Attention
    layers served by this backend load per-head k_scale / v_scale vectors (kv_scales.PerHeadKVCacheMethod) where vLLM's
    BaseKVCacheMethod insists on a float ($VLLM/model_executor/layers/quantization/kv_cache.py:103-106).  Done by patching at
    import like the reference patches vLLM (vllm_omni/patch.py:52-67): no edit of the reference tree.  Idempotent; returns
    False when vllm_omni is not importable (plain vLLM deployments keep vLLM's per-tensor method)."""
    try:
        from vllm_omni.quantization.component_config import ComponentQuantizationConfig
    except Exception:  # noqa: BLE001
        return False
    orig = ComponentQuantizationConfig.get_quant_method
    if getattr(orig, "_b200_per_head_kv", False):
        return True

    def get_quant_method(self, layer, prefix):
        delegated = orig(self, layer, prefix)
        ours = _per_head_method_for(layer, delegated)
        return delegated if ours is None else ours

    get_quant_method._b200_per_head_kv = True
    get_quant_method.__wrapped__ = orig
    ComponentQuantizationConfig.get_quant_method = get_quant_method
    return True


def register_omni_platform() -> None:
    """`vllm_omni.platform_plugins` hook: nothing to swap.  CudaOmniPlatform picks worker classes, stage configs and the
    diffusion attention backend only (vllm_omni/platforms/cuda/platform.py:24-81); the AR attention backend is chosen by
    vLLM's own selector ($VLLM/v1/attention/selector.py:111-117), which register() feeds.  Returns None = "no platform"."""
    return None
