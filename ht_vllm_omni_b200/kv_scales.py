"""Per-head FP8 (e4m3) KV-cache scale handling for the B200 backend.

The reference has no KV-scale code of its own: `vllm_omni/quantization` routes `get_quant_method` to vLLM
(vllm_omni/quantization/component_config.py:53-57, factory.py:52-73), and vLLM's `BaseKVCacheMethod` accepts only
PER-TENSOR scales ($VLLM/model_executor/layers/quantization/kv_cache.py:103-106 raises for anything but a float).  The
north star asks for per-head scales, so this module adds what is missing, keeping vLLM's attribute contract
(`layer._k_scale/_v_scale` fp32 device tensors read by the backend, `layer._k_scale_float/_v_scale_float` host floats,
$VLLM/model_executor/layers/attention/attention.py:94-119):

  * `resolve_checkpoint_scales`  the k_scale / v_scale / single-kv_scale / missing (-1) precedence rules of
                                 BaseKVCacheMethod.process_weights_after_loading (kv_cache.py:73-101), extended to
                                 vectors of Hkv values (per-head) and to TP head sharding;
  * `install_layer_scales`       writes them onto an attention layer (re-registering `_k_scale/_v_scale` as [Hkv]
                                 buffers when per-head);
  * `calc_kv_scales`             the dynamic first-step calibration (`absmax / range`, attention.py:531-539) per tensor
                                 or per KV head;
  * `PerHeadKVCacheMethod`       a BaseKVCacheMethod-shaped object (create_weights / process_weights_after_loading) that
                                 `ComponentQuantizationConfig.get_quant_method` can hand out for Attention layers.

Scale semantics are those of the kernels (include/b200attn.h): cache byte = e4m3_satfinite(x / scale[h]); the decode
kernel folds k_scale[h] into the logit scale and multiplies the output by v_scale[h].
"""
from __future__ import annotations

import torch

FP8_E4M3_MAX = 448.0
# vLLM's calibration ranges (envs.K_SCALE_CONSTANT / V_SCALE_CONSTANT defaults): scale = absmax / range
K_RANGE_DEFAULT = 200.0
V_RANGE_DEFAULT = 100.0


def _as_vector(x, num_kv_heads: int, name: str) -> torch.Tensor:
    t = torch.as_tensor(x, dtype=torch.float32).detach().cpu().reshape(-1)
    if t.numel() not in (1, num_kv_heads):
        raise ValueError(f"{name} must hold 1 (per-tensor) or {num_kv_heads} (per-head) values, got {t.numel()}")
    return t


def resolve_checkpoint_scales(k_scale, v_scale, num_kv_heads: int):
    """Checkpoint values -> (k, v) fp32 CPU vectors with 1 or Hkv elements each.

    Follows kv_cache.py:73-101: both valid (> 0) -> use as is; both invalid (< 0, nothing loaded) -> 1.0; exactly one
    valid (a lone `kv_scale` remapped onto k_scale) -> duplicate it.  Validity is judged per tensor (all elements).
    If one side is per-head and the other per-tensor, the per-tensor side is broadcast so both have the same length.
    """
    k = _as_vector(k_scale, num_kv_heads, "k_scale")
    v = _as_vector(v_scale, num_kv_heads, "v_scale")
    k_ok, v_ok = bool((k > 0).all()), bool((v > 0).all())
    if not k_ok and not v_ok:
        return torch.ones(1), torch.ones(1)
    if k_ok != v_ok:
        src = k if k_ok else v
        k, v = src.clone(), src.clone()
    n = max(k.numel(), v.numel())
    if k.numel() != n:
        k = k.expand(n).clone()
    if v.numel() != n:
        v = v.expand(n).clone()
    return k.contiguous(), v.contiguous()


def shard_scales(scale: torch.Tensor, total_num_kv_heads: int, tp_size: int, tp_rank: int) -> torch.Tensor:
    """Per-head scales of the whole model -> the slice this TP rank owns (vLLM head partitioning,
    $VLLM/model_executor/models/qwen3.py:78-95: Hkv/tp heads per rank, or one replicated head when tp > Hkv)."""
    if scale.numel() == 1:
        return scale
    if scale.numel() != total_num_kv_heads:
        raise ValueError(f"expected {total_num_kv_heads} per-head scales, got {scale.numel()}")
    if tp_size <= total_num_kv_heads:
        if total_num_kv_heads % tp_size:
            raise ValueError("num_kv_heads must be divisible by tp_size")
        per = total_num_kv_heads // tp_size
        return scale[tp_rank * per:(tp_rank + 1) * per].contiguous()
    if tp_size % total_num_kv_heads:
        raise ValueError("tp_size must be a multiple of num_kv_heads when it exceeds it")
    return scale[tp_rank // (tp_size // total_num_kv_heads)].reshape(1).contiguous()


def install_layer_scales(layer: torch.nn.Module, k_scale: torch.Tensor, v_scale: torch.Tensor) -> None:
    """Put (k, v) scales on an attention layer the way backends read them.  Per-tensor: copy into the existing 0-d
    buffers.  Per-head: replace the buffers by [Hkv] fp32 tensors on the same device (the kernels take a stride-1
    vector); the host floats then hold the MAX over heads (what a per-tensor consumer would need to stay in range)."""
    dev = layer._k_scale.device
    k = k_scale.to(torch.float32).reshape(-1)
    v = v_scale.to(torch.float32).reshape(-1)
    if k.numel() == 1 and v.numel() == 1:
        layer._k_scale.copy_(k.reshape(layer._k_scale.shape))
        layer._v_scale.copy_(v.reshape(layer._v_scale.shape))
    else:
        for name, val in (("_k_scale", k), ("_v_scale", v)):
            if name in getattr(layer, "_buffers", {}):
                del layer._buffers[name]
            elif hasattr(layer, name):
                delattr(layer, name)
            layer.register_buffer(name, val.to(dev).contiguous())
    layer._k_scale_float = float(k.max())
    layer._v_scale_float = float(v.max())


def calc_kv_scales(key: torch.Tensor, value: torch.Tensor, per_head: bool, k_range: float = K_RANGE_DEFAULT,
                   v_range: float = V_RANGE_DEFAULT, floor: float = 1e-8):
    """Dynamic calibration from one batch: scale = absmax / range (attention.py:531-536), per tensor or per KV head.
    key/value: [T, Hkv, D].  Returns fp32 tensors of 1 or Hkv elements on the input device (no host sync)."""
    if key.dim() != 3 or value.shape[:2] != key.shape[:2]:
        raise ValueError("key/value must be [T, Hkv, D]")
    if per_head:
        k = key.abs().amax(dim=(0, 2)).to(torch.float32) / k_range
        v = value.abs().amax(dim=(0, 2)).to(torch.float32) / v_range
    else:
        k = (key.abs().max().to(torch.float32) / k_range).reshape(1)
        v = (value.abs().max().to(torch.float32) / v_range).reshape(1)
    return k.clamp_min(floor), v.clamp_min(floor)


class PerHeadKVCacheMethod:
    """Quant method for Attention layers with an FP8 KV cache whose checkpoint carries per-head (or per-tensor) scales.

    Same life cycle as vLLM's BaseKVCacheMethod (kv_cache.py:18-172): `create_weights` adds loadable `k_scale/v_scale`
    parameters initialised to -1 (= "not in the checkpoint"), `process_weights_after_loading` turns them into the
    backend-facing `_k_scale/_v_scale` and deletes the parameters.  Differences: the parameters are [Hkv] vectors, the
    per-tensor ValueError is gone, and TP sharding of the vector is applied here.
    """

    def __init__(self, num_kv_heads: int, total_num_kv_heads: int | None = None, tp_size: int = 1, tp_rank: int = 0):
        self.num_kv_heads = num_kv_heads
        self.total_num_kv_heads = total_num_kv_heads or num_kv_heads
        self.tp_size, self.tp_rank = tp_size, tp_rank

    def create_weights(self, layer: torch.nn.Module) -> None:
        n = self.total_num_kv_heads
        layer.k_scale = torch.nn.Parameter(torch.full((n,), -1.0), requires_grad=False)
        layer.v_scale = torch.nn.Parameter(torch.full((n,), -1.0), requires_grad=False)

    def apply(self, layer: torch.nn.Module):
        raise RuntimeError(f"{self.__class__.__name__}.apply should not be called.")

    def process_weights_after_loading(self, layer: torch.nn.Module) -> None:
        if not hasattr(layer, "k_scale"):
            return   # nothing to process (weight reloading)
        fp8 = getattr(layer, "kv_cache_dtype", "auto") in ("fp8", "fp8_e4m3")
        if fp8 and not getattr(layer, "calculate_kv_scales", False):
            k, v = resolve_checkpoint_scales(layer.k_scale.data, layer.v_scale.data, self.total_num_kv_heads)
            if k.numel() > 1 and bool((k == k[0]).all()) and bool((v == v[0]).all()):
                k, v = k[:1], v[:1]   # a uniform vector is a per-tensor scale
            k = shard_scales(k, self.total_num_kv_heads, self.tp_size, self.tp_rank)
            v = shard_scales(v, self.total_num_kv_heads, self.tp_size, self.tp_rank)
            install_layer_scales(layer, k, v)
        del layer.k_scale
        del layer.v_scale
