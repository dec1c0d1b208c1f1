"""ht_vllm_omni_b200 — B200-native paged-decode attention for the AR stages of ht-vllm-omni.

One hot path, nothing else: the KV write (+ FP8 quantisation) and split-KV paged decode attention that the
Qwen3-TTS talker / Qwen3-Omni thinker stages run once per layer per step, as hand-written sm_100a CUDA behind
a C ABI (``libb200attn.so``), surfaced to vLLM through its AttentionBackend plugin API.

    ops      torch.library custom ops over the C ABI (``torch.ops.b200attn.*``)
    backend  vLLM AttentionBackend / AttentionMetadataBuilder / AttentionImpl
    plugin   ``register()`` entry point for ``vllm_omni.general_plugins`` / ``vllm.general_plugins``
    kv_scales  per-head FP8 KV scale handling (what vllm_omni/quantization lacks)
    parallel KV-head sharding across ranks + output all-gather
"""
__version__ = "0.1.0"

from ._lib import DT_BF16, DT_FP16, DT_FP8_E4M3, B200AttnError, lib_path, load_library  # noqa: F401
