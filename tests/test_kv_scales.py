"""Per-head FP8 KV scale handling (ht_vllm_omni_b200/kv_scales.py): checkpoint precedence rules of vLLM's
BaseKVCacheMethod ($VLLM/model_executor/layers/quantization/kv_cache.py:73-101) extended to per-head vectors, TP
sharding of the vector, dynamic calibration, and the layer attribute contract the backend reads."""
import numpy as np
import pytest
import torch

from ht_vllm_omni_b200 import kv_scales as S
from oracle import oracle as O


class Layer(torch.nn.Module):
    def __init__(self, kv_cache_dtype="fp8"):
        super().__init__()
        self.register_buffer("_k_scale", torch.tensor(1.0))
        self.register_buffer("_v_scale", torch.tensor(1.0))
        self._k_scale_float = self._v_scale_float = 1.0
        self.kv_cache_dtype = kv_cache_dtype
        self.calculate_kv_scales = False


def test_checkpoint_precedence_rules():
    k, v = S.resolve_checkpoint_scales(torch.tensor(0.5), torch.tensor(0.25), 4)
    assert k.tolist() == [0.5] and v.tolist() == [0.25]
    k, v = S.resolve_checkpoint_scales(torch.tensor(-1.0), torch.tensor(-1.0), 4)       # nothing in the checkpoint
    assert k.tolist() == [1.0] and v.tolist() == [1.0]
    k, v = S.resolve_checkpoint_scales(torch.tensor(0.5), torch.tensor(-1.0), 4)        # lone kv_scale -> duplicated
    assert k.tolist() == [0.5] and v.tolist() == [0.5]
    k, v = S.resolve_checkpoint_scales(torch.tensor([0.1, 0.2, 0.3, 0.4]), torch.tensor(0.5), 4)   # mixed -> broadcast
    assert k.numel() == 4 and v.tolist() == [0.5] * 4
    with pytest.raises(ValueError):
        S.resolve_checkpoint_scales(torch.ones(3), torch.ones(1), 4)


def test_tp_sharding_of_per_head_scales():
    s = torch.tensor([0.1, 0.2, 0.3, 0.4])
    assert S.shard_scales(s, 4, 2, 1).tolist() == pytest.approx([0.3, 0.4])
    assert S.shard_scales(s, 4, 4, 2).tolist() == pytest.approx([0.3])
    assert S.shard_scales(s, 4, 8, 5).tolist() == pytest.approx([0.3])    # tp > Hkv: head 5 // 2 replicated
    assert S.shard_scales(torch.tensor([0.7]), 4, 8, 5).tolist() == pytest.approx([0.7])
    with pytest.raises(ValueError):
        S.shard_scales(s, 4, 3, 0)


def test_quant_method_life_cycle_per_head_and_per_tensor():
    layer = Layer()
    m = S.PerHeadKVCacheMethod(num_kv_heads=2, total_num_kv_heads=4, tp_size=2, tp_rank=1)
    m.create_weights(layer)
    assert layer.k_scale.shape == (4,) and float(layer.k_scale[0]) == -1.0
    layer.k_scale.data.copy_(torch.tensor([0.1, 0.2, 0.3, 0.4]))
    layer.v_scale.data.copy_(torch.tensor([0.5, 0.6, 0.7, 0.8]))
    m.process_weights_after_loading(layer)
    assert not hasattr(layer, "k_scale") and not hasattr(layer, "v_scale")
    assert layer._k_scale.tolist() == pytest.approx([0.3, 0.4]) and layer._v_scale.tolist() == pytest.approx([0.7, 0.8])
    assert layer._k_scale.dtype == torch.float32 and layer._k_scale.is_contiguous()
    assert layer._k_scale_float == pytest.approx(0.4) and layer._v_scale_float == pytest.approx(0.8)
    assert "_k_scale" in dict(layer.named_buffers())           # still moves with model.to(device)

    layer = Layer()                                            # uniform vector collapses to vLLM's per-tensor case
    m = S.PerHeadKVCacheMethod(4)
    m.create_weights(layer)
    layer.k_scale.data.fill_(0.25)
    layer.v_scale.data.fill_(0.5)
    m.process_weights_after_loading(layer)
    assert layer._k_scale.shape == () and float(layer._k_scale) == 0.25 and layer._v_scale_float == 0.5

    layer = Layer()                                            # missing scales -> 1.0 (kv_cache.py:85-89)
    m.create_weights(layer)
    m.process_weights_after_loading(layer)
    assert float(layer._k_scale) == 1.0 and float(layer._v_scale) == 1.0
    m.process_weights_after_loading(layer)                     # reload: nothing left to process, no error

    layer = Layer("auto")                                      # unquantised cache: scales untouched
    m.create_weights(layer)
    layer.k_scale.data.fill_(0.25)
    m.process_weights_after_loading(layer)
    assert float(layer._k_scale) == 1.0


def test_dynamic_calibration_keeps_values_in_range_and_matches_write_semantics():
    g = torch.Generator().manual_seed(0)
    key = torch.randn(50, 4, 128, generator=g) * torch.tensor([1.0, 10.0, 1e-4, 3.0]).view(1, 4, 1)
    val = torch.randn(50, 4, 128, generator=g)
    k, v = S.calc_kv_scales(key, val, per_head=True)
    assert k.shape == (4,) and v.shape == (4,)
    assert torch.allclose(k, key.abs().amax(dim=(0, 2)) / S.K_RANGE_DEFAULT)
    kt, vt = S.calc_kv_scales(key, val, per_head=False)
    assert kt.shape == (1,) and float(kt) == pytest.approx(float(key.abs().max()) / S.K_RANGE_DEFAULT)
    # x / scale stays inside e4m3's finite range, and per-head scales quantise the small-magnitude head far better
    assert float((key / k.view(1, 4, 1)).abs().max()) <= S.K_RANGE_DEFAULT * (1 + 1e-6) < S.FP8_E4M3_MAX

    def roundtrip_err(scale):
        sc = scale.numpy() if scale.numel() > 1 else np.repeat(scale.numpy(), 4)
        x = key.numpy().astype(np.float32)
        deq = O.e4m3_decode(O.e4m3_encode_satfinite(x / sc[None, :, None])) * sc[None, :, None]
        return np.abs(deq - x)[:, 2].mean()      # head 2 has 1e-4 of head 1's magnitude: e4m3 subnormals under a shared scale

    assert roundtrip_err(k) < 0.2 * roundtrip_err(kt)


def test_install_layer_scales_per_tensor_in_place():
    layer = Layer()
    before = layer._k_scale.data_ptr()
    S.install_layer_scales(layer, torch.tensor([0.3]), torch.tensor([0.6]))
    assert layer._k_scale.data_ptr() == before and float(layer._k_scale) == pytest.approx(0.3)
    assert layer._v_scale_float == pytest.approx(0.6)


def test_plugin_installs_per_head_method_into_the_reference_quant_routing(monkeypatch):
    """plugin.register() wraps ComponentQuantizationConfig.get_quant_method (vllm_omni/quantization/component_config.py:53-57)
    without editing the reference: Attention layers served by this backend with an FP8 cache get PerHeadKVCacheMethod where
    vLLM would have handed out its per-tensor BaseKVCacheMethod; everything else is passed through untouched.  vllm_omni is
    not importable in this image, so a stand-in module with the reference's class shape is installed."""
    import sys
    import types

    from vllm.model_executor.layers.quantization.kv_cache import BaseKVCacheMethod

    from ht_vllm_omni_b200 import backend as BK
    from ht_vllm_omni_b200 import plugin

    class ComponentQuantizationConfig:           # shape of the reference class: resolve(prefix) -> config -> get_quant_method
        def __init__(self, method):
            self.method = method

        def get_quant_method(self, layer, prefix):
            return self.method

    mod = types.ModuleType("vllm_omni.quantization.component_config")
    mod.ComponentQuantizationConfig = ComponentQuantizationConfig
    for name in ("vllm_omni", "vllm_omni.quantization"):
        monkeypatch.setitem(sys.modules, name, types.ModuleType(name))
    monkeypatch.setitem(sys.modules, "vllm_omni.quantization.component_config", mod)

    assert plugin.install_per_head_kv_scales() is True
    assert plugin.install_per_head_kv_scales() is True            # idempotent: wrapped once
    assert ComponentQuantizationConfig.get_quant_method.__wrapped__.__name__ == "get_quant_method"
    assert not hasattr(ComponentQuantizationConfig.get_quant_method.__wrapped__, "__wrapped__")

    per_tensor = BaseKVCacheMethod.__new__(BaseKVCacheMethod)
    ours = Layer("fp8"); ours.attn_backend = BK.B200PagedAttentionBackend; ours.num_kv_heads = 4
    m = ComponentQuantizationConfig(per_tensor).get_quant_method(ours, "talker.layers.0.self_attn.attn")
    assert isinstance(m, S.PerHeadKVCacheMethod) and m.num_kv_heads == 4 and m.total_num_kv_heads == 4
    # ... and it loads a per-head checkpoint vector onto the layer
    m.create_weights(ours)
    ours.k_scale.data.copy_(torch.tensor([0.1, 0.2, 0.3, 0.4])); ours.v_scale.data.copy_(torch.tensor([0.5, 0.6, 0.7, 0.8]))
    m.process_weights_after_loading(ours)
    assert ours._k_scale.tolist() == pytest.approx([0.1, 0.2, 0.3, 0.4]) and ours._v_scale.tolist() == pytest.approx([0.5, 0.6, 0.7, 0.8])

    other_backend = Layer("fp8"); other_backend.attn_backend = type("Other", (), {}); other_backend.num_kv_heads = 4
    assert ComponentQuantizationConfig(per_tensor).get_quant_method(other_backend, "x") is per_tensor
    bf16 = Layer("auto"); bf16.attn_backend = BK.B200PagedAttentionBackend; bf16.num_kv_heads = 4
    assert ComponentQuantizationConfig(per_tensor).get_quant_method(bf16, "x") is per_tensor
    assert ComponentQuantizationConfig(None).get_quant_method(ours, "x") is None          # no config for the prefix: unchanged
    linear_method = object()
    assert ComponentQuantizationConfig(linear_method).get_quant_method(ours, "x") is linear_method
