"""Host-side checks of the vLLM plugin surface (no GPU): the class answers vLLM's selector / KV allocator ask for,
registration, per-token row expansion, and the hooks' argument handling.  Mirrors how the reference tests its
runners: dummy objects standing in for vLLM's config / layer ($REF/tests/worker/*)."""
import types

import numpy as np
import pytest
import torch

vllm_backend = pytest.importorskip("vllm.v1.attention.backend")

from ht_vllm_omni_b200 import backend as B  # noqa: E402
from ht_vllm_omni_b200 import plugin  # noqa: E402


def test_backend_static_surface():
    be = B.B200PagedAttentionBackend
    assert issubclass(be, vllm_backend.AttentionBackend)
    assert be.get_name() == "CUSTOM"
    assert be.get_impl_cls() is B.B200PagedAttentionImpl and be.get_builder_cls() is B.B200PagedAttentionMetadataBuilder
    assert be.forward_includes_kv_cache_update is True and be.accept_output_buffer is True   # B200ATTN_SEPARATE_KV_UPDATE unset
    bc = be.get_builder_cls()
    assert bc.get_cudagraph_support(None, None) == vllm_backend.AttentionCGSupport.UNIFORM_BATCH
    assert be.get_kv_cache_shape(10, 16, 4, 128, "fp8") == (10, 2, 16, 4, 128)
    assert be.get_kv_cache_block_dim(16, 4, 128) == 0
    assert be.get_kv_cache_stride_order() == (0, 1, 3, 2, 4)
    assert sorted(be.get_kv_cache_stride_order(True)) == list(range(6))
    assert be.get_required_kv_cache_layout() == "HND"
    assert be.supports_block_size(16) and be.supports_block_size(128) and not be.supports_block_size(8)
    assert be.supports_head_size(128) and be.supports_head_size(64) and be.supports_head_size(256) and not be.supports_head_size(96)
    assert be.supports_kv_cache_dtype("fp8") and be.supports_kv_cache_dtype("fp8_e4m3") and not be.supports_kv_cache_dtype("fp8_e5m2")
    assert be.supports_per_head_quant_scales()
    with pytest.raises(ValueError):
        be.get_kv_cache_shape(10, 8, 4, 128)


def _validate(**over):
    cap = types.SimpleNamespace(major=10, minor=0)
    kw = dict(head_size=128, dtype=torch.bfloat16, kv_cache_dtype="fp8", block_size=16, use_mla=False, has_sink=False,
              use_sparse=False, use_mm_prefix=False, use_per_head_quant_scales=True, device_capability=cap,
              attn_type=vllm_backend.AttentionType.DECODER)
    kw.update(over)
    return B.B200PagedAttentionBackend.validate_configuration(**kw)


def test_validate_configuration_returns_reasons_not_exceptions():
    assert _validate() == []
    assert _validate(kv_cache_dtype="bfloat16") == []
    assert any("head_size" in r for r in _validate(head_size=96))
    assert any("kv_cache_dtype" in r for r in _validate(kv_cache_dtype="fp8_e5m2"))
    assert any("block_size" in r for r in _validate(block_size=8))
    assert any("compute capability" in r for r in _validate(device_capability=types.SimpleNamespace(major=9, minor=0)))
    assert any("MLA" in r for r in _validate(use_mla=True))
    assert any("sink" in r for r in _validate(has_sink=True))
    assert any("must match" in r for r in _validate(kv_cache_dtype="float16"))


def test_kv_cache_allocation_gives_hnd_pages():
    """Replay _reshape_kv_cache_tensors ($VLLM/v1/worker/gpu_model_runner.py:6891-6984): raw buffer -> physical shape ->
    permute back to the logical shape.  One (block, K|V, head) page must then be block_size*D contiguous elements."""
    be = B.B200PagedAttentionBackend
    nb, bs, H, D = 6, 16, 4, 128
    shape = be.get_kv_cache_shape(nb, bs, H, D, "fp8")
    order = be.get_kv_cache_stride_order()
    phys = tuple(shape[i] for i in order)
    inv = [order.index(i) for i in range(len(order))]
    raw = torch.zeros(int(np.prod(shape)), dtype=torch.int8)
    cache = raw.view(torch.uint8).view(phys).permute(*inv)
    assert tuple(cache.shape) == shape
    k = cache.unbind(1)[0]                       # [nb, bs, H, D]
    assert k.stride() == (2 * H * bs * D, D, bs * D, 1)


def test_plugin_register_is_idempotent():
    from vllm.v1.attention.backends.registry import _ATTN_OVERRIDES, AttentionBackendEnum

    assert plugin.register() == "CUSTOM"
    assert plugin.register() == "CUSTOM"
    assert _ATTN_OVERRIDES[AttentionBackendEnum.CUSTOM] == B.BACKEND_PATH
    assert AttentionBackendEnum.CUSTOM.get_class() is B.B200PagedAttentionBackend
    assert plugin.register_omni_platform() is None


def test_expand_query_rows_matches_definition():
    # requests: q_len 1 (ctx 9), q_len 3 (ctx 20 -> seq 23), q_len 2 (seq 2, fresh prefill), + 2 padding tokens
    qsl = torch.tensor([0, 1, 4, 6], dtype=torch.int32)
    seq = torch.tensor([10, 23, 2], dtype=torch.int32)
    lens, rmap = B.expand_query_rows(qsl, seq, 8)
    assert lens.dtype == torch.int32 and rmap.dtype == torch.int32
    assert lens.tolist() == [10, 21, 22, 23, 1, 2, 0, 0]
    assert rmap.tolist()[:6] == [0, 1, 1, 1, 2, 2]
    # pure decode degenerates to the identity
    qsl = torch.arange(0, 6, dtype=torch.int32)
    seq = torch.tensor([5, 1, 7, 3, 9], dtype=torch.int32)
    lens, rmap = B.expand_query_rows(qsl, seq, 5)
    assert lens.tolist() == seq.tolist() and rmap.tolist() == [0, 1, 2, 3, 4]


def test_impl_rejects_unsupported_features_at_construction():
    Impl = B.B200PagedAttentionImpl
    ok = Impl(16, 128, 128 ** -0.5, 2, None, None, "fp8")
    assert ok.num_queries_per_kv == 8 and ok.kv_cache_dtype == "fp8"
    for kw in (dict(alibi_slopes=[1.0] * 16), dict(sliding_window=128), dict(logits_soft_cap=30.0),
               dict(kv_cache_dtype="fp8_e5m2"), dict(attn_type=vllm_backend.AttentionType.ENCODER_ONLY)):
        args = dict(num_heads=16, head_size=128, scale=0.1, num_kv_heads=2)
        args.update(kw)
        with pytest.raises(NotImplementedError):
            Impl(**args)
    with pytest.raises(NotImplementedError):
        Impl(16, 96, 0.1, 2)


def test_impl_profile_run_and_no_cpu_fallback():
    impl = B.B200PagedAttentionImpl(16, 128, 128 ** -0.5, 2, None, None, "fp8")
    out = torch.ones(4, 16 * 128)
    assert impl.forward(None, torch.zeros(4, 16 * 128), None, None, torch.zeros(0), None, out) is out
    assert float(out.abs().sum()) == 0.0                      # attn_metadata None = profiling run: zero-fill
    layer = types.SimpleNamespace(_k_scale=torch.ones(1), _v_scale=torch.ones(1))
    cache = torch.zeros(4, 2, 16, 2, 128, dtype=torch.uint8)
    key = torch.zeros(3, 2 * 128, dtype=torch.bfloat16)
    with pytest.raises(Exception):                            # CPU tensors: the CUDA op refuses, nothing falls back
        impl.do_kv_cache_update(layer, key, key, cache, torch.zeros(3, dtype=torch.int64))
    with pytest.raises(B.B200AttnError):
        impl._planes(torch.zeros(2, 4, 16, 2, 128, dtype=torch.uint8))


def test_aligned_grid_policy():
    """ops.aligned_grid: pair-aligned decode grids only when every pair gets >= 2 CTAs and >= 80 % of the slots stay used."""
    from ht_vllm_omni_b200.ops import aligned_grid
    assert aligned_grid(296, 64, 16, 2) == 256      # c2: 128 pairs x 2
    assert aligned_grid(296, 32, 32, 4) == 256      # thinker B=32: 128 pairs x 2
    assert aligned_grid(296, 32, 16, 2) == 256      # 64 pairs x 4
    assert aligned_grid(296, 32, 8, 1) == 256       # thinker at tp 4: 32 pairs x 8 (9 would cut clusters of 2 across pairs)
    assert aligned_grid(296, 1, 8, 1) == 296        # one pair: 296 CTAs (already even)
    assert aligned_grid(296, 1, 16, 2) == 296       # 2 pairs x 148
    assert aligned_grid(296, 100, 16, 2) == 296     # 200 pairs: k = 1 -> default
    assert aligned_grid(296, 256, 16, 2) == 296     # more pairs than CTAs
    assert aligned_grid(296, 50, 16, 2) == 296      # 100 pairs x 2 = 200 < 80 % of 296 -> default
    assert aligned_grid(296, 0, 16, 2) == 296
    for b in range(1, 300):
        g = aligned_grid(296, b, 16, 2)
        assert 0.8 * 296 <= g <= 296
