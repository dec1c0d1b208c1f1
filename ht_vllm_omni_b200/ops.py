"""torch.library custom ops over the C ABI: ``torch.ops.b200attn.{kv_write, decode_plan, paged_decode, merge_lse}``.

They are what ``backend.B200PagedAttentionImpl`` calls from ``do_kv_cache_update`` / ``forward`` — the same two
call sites the reference reaches through ``torch.ops.vllm.unified_kv_cache_update`` and
``unified_attention_with_output`` ($VLLM/model_executor/layers/attention/attention.py:691-761).  Registered as
custom ops (with fake impls) so they trace under torch.compile and capture into CUDA graphs like the ops they
replace.  PyTorch only supplies device memory and the current stream here; all arithmetic is in libb200attn.

Tensor conventions (same as vLLM's): key/value/query ``[T, H, D]``; ``k_cache``/``v_cache`` are the two planes of
the paged cache viewed as ``[num_blocks, block_size, Hkv, D]`` with ANY strides (HND physical layout preferred);
``slot_mapping`` int64 ``[T]`` (-1 = padding); ``block_table`` int32 ``[B, max_blocks]``; ``seq_lens`` int32 ``[B]``;
``k_scale``/``v_scale`` fp32 with 1 element (per-tensor) or Hkv elements (per-head).
"""
from __future__ import annotations

import torch

from . import _lib
from ._lib import DECODE_FUSED_MERGE, DT_BF16, DT_FP16, DT_FP8_E4M3, B200AttnError, check, load_library

_FP8_DTYPES = (torch.uint8, torch.float8_e4m3fn)


def _dt(t: torch.Tensor) -> int:
    if t.dtype == torch.bfloat16:
        return DT_BF16
    if t.dtype == torch.float16:
        return DT_FP16
    if t.dtype in _FP8_DTYPES:
        return DT_FP8_E4M3
    raise B200AttnError(f"unsupported dtype {t.dtype}")


def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def _require_cuda(*ts: torch.Tensor) -> None:
    for t in ts:
        if t is not None and not t.is_cuda:
            raise B200AttnError("b200attn ops need CUDA tensors: there is no CPU fallback for this path")


def _scale_stride(scale: torch.Tensor, num_kv_heads: int, other: torch.Tensor | None = None) -> int:
    """Element stride of the per-tensor (0) / per-head (1) scale vectors.  The kernels apply ONE stride to k_scale and
    v_scale, so `other` (the second scale) must have the same number of elements."""
    n = scale.numel()
    if other is not None and other.numel() != n:
        raise B200AttnError(f"k_scale and v_scale must have the same number of elements ({n} vs {other.numel()})")
    if scale.dtype != torch.float32 or (other is not None and other.dtype != torch.float32):
        raise B200AttnError("kv scales must be fp32")
    if n == 1:
        return 0
    if n == num_kv_heads and scale.is_contiguous() and (other is None or other.is_contiguous()):
        return 1
    raise B200AttnError(f"kv scale must have 1 or Hkv={num_kv_heads} contiguous elements, got shape {tuple(scale.shape)}")


def _cache_strides(k_cache: torch.Tensor, v_cache: torch.Tensor):
    if k_cache.dim() != 4 or v_cache.shape != k_cache.shape or v_cache.stride() != k_cache.stride():
        raise B200AttnError("k_cache / v_cache must be [num_blocks, block_size, Hkv, D] views with identical strides")
    if k_cache.stride(3) != 1:
        raise B200AttnError("cache head_dim must be contiguous")
    return k_cache.stride(0), k_cache.stride(1), k_cache.stride(2)


def default_ctas(kv_dtype: torch.dtype) -> int:
    """Persistent decode grid for this device: 2 CTAs/SM."""
    n = load_library().b200attn_default_ctas(DT_FP8_E4M3 if kv_dtype in _FP8_DTYPES else DT_BF16)
    if n <= 0:
        raise B200AttnError("no CUDA device: libb200attn has no CPU fallback")
    return n


def plan_numel(max_num_seqs: int) -> int:
    return load_library().b200attn_plan_bytes(max_num_seqs) // 4


def workspace_numel(head_dim: int, num_ctas: int) -> int:
    return load_library().b200attn_workspace_bytes(head_dim, num_ctas) // 4


# ------------------------------------------------------------------------------------------------
# kv_write
# ------------------------------------------------------------------------------------------------
@torch.library.custom_op("b200attn::kv_write", mutates_args=("k_cache", "v_cache"), device_types="cuda")
def kv_write(key: torch.Tensor, value: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
             slot_mapping: torch.Tensor, k_scale: torch.Tensor, v_scale: torch.Tensor) -> None:
    _require_cuda(key, value, k_cache, v_cache, slot_mapping)
    lib = load_library()
    T = slot_mapping.shape[0]
    if T == 0:
        return
    if key.dim() != 3 or value.shape != key.shape or key.stride(2) != 1 or value.stride(2) != 1:
        raise B200AttnError("key/value must be [T, Hkv, D] with contiguous head_dim")
    if slot_mapping.dtype != torch.int64 or not slot_mapping.is_contiguous():
        raise B200AttnError("slot_mapping must be contiguous int64")
    Hkv, D = key.shape[1], key.shape[2]
    sb, ss, sh = _cache_strides(k_cache, v_cache)
    kvdt = _dt(k_cache)
    sstride = _scale_stride(k_scale, Hkv, v_scale) if kvdt == DT_FP8_E4M3 else 0
    check(lib.b200attn_kv_write(
        key.data_ptr(), value.data_ptr(), key.stride(0), key.stride(1), value.stride(0), value.stride(1),
        k_cache.data_ptr(), v_cache.data_ptr(), sb, ss, sh, slot_mapping.data_ptr(), T, Hkv, D, k_cache.shape[1],
        _dt(key), kvdt, k_scale.data_ptr(), v_scale.data_ptr(), sstride, _stream(key)), "b200attn_kv_write")


@kv_write.register_fake
def _(key, value, k_cache, v_cache, slot_mapping, k_scale, v_scale) -> None:
    return None


# ------------------------------------------------------------------------------------------------
# fp8_per_token_head: dynamic per-(token, head) scales in float32 side caches
# ------------------------------------------------------------------------------------------------
def per_token_head_views(kv_cache: torch.Tensor, num_kv_heads: int, head_dim: int):
    """The four views of a per-token-head cache allocation.  vLLM budgets, per block, 2*bs*Hkv*D bytes of e4m3 data plus
    2*bs*Hkv float32 scales ($VLLM/v1/kv_cache_interface.py:150-163) and hands the backend `[nb, 2, bs, Hkv, D + 4]` bytes
    (the shape TritonAttentionBackend.get_kv_cache_shape declares, triton_attn.py:312-334).  This backend lays a block out as
        [K data: Hkv][bs][D] | [V data: Hkv][bs][D] | [K scales: Hkv][bs] f32 | [V scales: Hkv][bs] f32
    so the data planes are the same HND pages as every other cache dtype (TMA boxes of 16 tokens x 128 bytes) and 16
    consecutive tokens' scales are one 64-byte run.  Returns (k_cache, v_cache) uint8 `[nb, bs, Hkv, D]` and
    (k_scale_cache, v_scale_cache) float32 `[nb, bs, Hkv]`, all views of `kv_cache`'s memory (any tensor whose blocks are
    `2*bs*Hkv*(D+4)` contiguous bytes each, first dim = blocks)."""
    if kv_cache.dim() != 5 or kv_cache.shape[1] != 2 or kv_cache.shape[3] != num_kv_heads or kv_cache.shape[4] != head_dim + 4 \
            or kv_cache.element_size() != 1:
        raise B200AttnError(f"per-token-head cache must be 1-byte [nb, 2, bs, Hkv, D + 4], got {tuple(kv_cache.shape)} {kv_cache.dtype}")
    nb, _, bs, H, _ = kv_cache.shape
    D = head_dim
    blk = 2 * bs * H * (D + 4)
    if kv_cache.stride(0) != blk:
        raise B200AttnError("per-token-head cache: every block must be one contiguous run of 2*bs*Hkv*(D+4) bytes")
    flat = torch.as_strided(kv_cache.view(torch.uint8), (nb * blk,), (1,), kv_cache.storage_offset())
    k = torch.as_strided(flat, (nb, bs, H, D), (blk, D, bs * D, 1), flat.storage_offset())
    v = torch.as_strided(flat, (nb, bs, H, D), (blk, D, bs * D, 1), flat.storage_offset() + H * bs * D)
    f32 = flat.view(torch.float32)
    so = 2 * H * bs * D // 4
    ks = torch.as_strided(f32, (nb, bs, H), (blk // 4, 1, bs), f32.storage_offset() + so)
    vs = torch.as_strided(f32, (nb, bs, H), (blk // 4, 1, bs), f32.storage_offset() + so + H * bs)
    return k, v, ks, vs


@torch.library.custom_op("b200attn::kv_write_per_token_head", mutates_args=("k_cache", "v_cache", "k_scale_cache", "v_scale_cache"),
                         device_types="cuda")
def kv_write_per_token_head(key: torch.Tensor, value: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                            k_scale_cache: torch.Tensor, v_scale_cache: torch.Tensor, slot_mapping: torch.Tensor,
                            ieee_div: bool = False) -> None:
    """KV write with one dynamic scale per (token, head): scale = max(absmax / 448, 1e-6) into the float32 side caches
    `[nb, bs, Hkv]` (any strides), bytes = e4m3(clamp(x / scale)).  `ieee_div`: IEEE divisions instead of the compiled
    reference's div.full.f32 (include/b200attn.h)."""
    _require_cuda(key, value, k_cache, v_cache, k_scale_cache, v_scale_cache, slot_mapping)
    T = slot_mapping.shape[0]
    if T == 0:
        return
    if key.dim() != 3 or value.shape != key.shape or key.stride(2) != 1 or value.stride(2) != 1:
        raise B200AttnError("key/value must be [T, Hkv, D] with contiguous head_dim")
    if slot_mapping.dtype != torch.int64 or not slot_mapping.is_contiguous():
        raise B200AttnError("slot_mapping must be contiguous int64")
    if _dt(k_cache) != DT_FP8_E4M3:
        raise B200AttnError("per-token-head scales need an fp8 (1-byte) cache")
    Hkv, D = key.shape[1], key.shape[2]
    sb, ss, sh = _cache_strides(k_cache, v_cache)
    if k_scale_cache.dtype != torch.float32 or v_scale_cache.dtype != torch.float32 or k_scale_cache.shape != v_scale_cache.shape \
            or k_scale_cache.stride() != v_scale_cache.stride() or tuple(k_scale_cache.shape) != (k_cache.shape[0], k_cache.shape[1], Hkv):
        raise B200AttnError("scale caches must be float32 [num_blocks, block_size, Hkv] views with identical strides")
    check(load_library().b200attn_kv_write_per_token_head(
        key.data_ptr(), value.data_ptr(), key.stride(0), key.stride(1), value.stride(0), value.stride(1), k_cache.data_ptr(),
        v_cache.data_ptr(), sb, ss, sh, k_scale_cache.data_ptr(), v_scale_cache.data_ptr(), k_scale_cache.stride(0),
        k_scale_cache.stride(1), k_scale_cache.stride(2), slot_mapping.data_ptr(), T, Hkv, D, k_cache.shape[1], _dt(key),
        1 if ieee_div else 0, _stream(key)), "b200attn_kv_write_per_token_head")


@kv_write_per_token_head.register_fake
def _(key, value, k_cache, v_cache, k_scale_cache, v_scale_cache, slot_mapping, ieee_div=False) -> None:
    return None


@torch.library.custom_op("b200attn::paged_decode_per_token_head", mutates_args=("out", "plan", "workspace", "lse"), device_types="cuda")
def paged_decode_per_token_head(query: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, k_scale_cache: torch.Tensor,
                                v_scale_cache: torch.Tensor, block_table: torch.Tensor, seq_lens: torch.Tensor, out: torch.Tensor,
                                plan: torch.Tensor, workspace: torch.Tensor, sm_scale: float, num_ctas: int, lse: torch.Tensor,
                                row_map: torch.Tensor, fresh_tokens: int = -1) -> None:
    """paged_decode over a cache with per-(token, head) scales (b200attn_decode_per_token_head); the scale caches must have
    slot stride 1 (`per_token_head_views`).  Split-KV partials are merged in-kernel."""
    _require_cuda(query, k_cache, v_cache, k_scale_cache, v_scale_cache, block_table, seq_lens, out, plan, workspace)
    if seq_lens.shape[0] == 0:
        return
    lse_ = lse if lse.numel() > 0 else None
    rm = row_map if row_map.numel() > 0 else None
    if query.dim() != 3 or out.dim() != 3 or query.stride(2) != 1 or out.stride(2) != 1 or query.dtype != out.dtype:
        raise B200AttnError("query/out must be [B, Hq, D] of the same dtype with contiguous head_dim")
    if block_table.dtype != torch.int32 or seq_lens.dtype != torch.int32 or block_table.stride(1) != 1:
        raise B200AttnError("block_table / seq_lens must be int32 (block_table rows contiguous)")
    if _dt(k_cache) != DT_FP8_E4M3 or k_scale_cache.dtype != torch.float32 or k_scale_cache.stride() != v_scale_cache.stride():
        raise B200AttnError("per-token-head decode needs an fp8 cache and float32 scale caches with identical strides")
    sb, ss, sh = _cache_strides(k_cache, v_cache)
    check(load_library().b200attn_decode_per_token_head(
        query.data_ptr(), query.stride(0), query.stride(1), out.data_ptr(), out.stride(0), out.stride(1), k_cache.data_ptr(),
        v_cache.data_ptr(), sb, ss, sh, k_cache.shape[0], k_scale_cache.data_ptr(), v_scale_cache.data_ptr(), k_scale_cache.stride(0),
        k_scale_cache.stride(1), k_scale_cache.stride(2), block_table.data_ptr(), block_table.stride(0), seq_lens.data_ptr(),
        seq_lens.shape[0], query.shape[1], k_cache.shape[2], query.shape[2], k_cache.shape[1], _dt(query), float(sm_scale),
        plan.data_ptr(), num_ctas, workspace.data_ptr(), workspace.numel() * 4, lse_.data_ptr() if lse_ is not None else None,
        rm.data_ptr() if rm is not None else None, int(fresh_tokens), DECODE_FUSED_MERGE, _stream(query)),
        "b200attn_decode_per_token_head")


@paged_decode_per_token_head.register_fake
def _(query, k_cache, v_cache, k_scale_cache, v_scale_cache, block_table, seq_lens, out, plan, workspace, sm_scale, num_ctas, lse,
      row_map, fresh_tokens=-1) -> None:
    return None


# ------------------------------------------------------------------------------------------------
# rope_kv_write: rotary embedding of q / k in place + KV write, one launch (do_rope_and_kv_cache_update)
# ------------------------------------------------------------------------------------------------
@torch.library.custom_op("b200attn::rope_kv_write", mutates_args=("query", "key", "k_cache", "v_cache"), device_types="cuda")
def rope_kv_write(query: torch.Tensor, key: torch.Tensor, value: torch.Tensor, positions: torch.Tensor,
                  cos_sin_cache: torch.Tensor, is_neox: bool, k_cache: torch.Tensor, v_cache: torch.Tensor,
                  slot_mapping: torch.Tensor, k_scale: torch.Tensor, v_scale: torch.Tensor, round_each_op: bool = False) -> None:
    """`round_each_op`: torch-native rounding (forward_native) instead of the compiled op's fp32 arithmetic.
    query [T, Hq, D], key / value [T, Hkv, D] (head_dim contiguous), positions int64 [T], cos_sin_cache
    [max_pos, rotary_dim] in the activation dtype (vLLM's RotaryEmbedding.cos_sin_cache after _match_cos_sin_cache_dtype)."""
    _require_cuda(query, key, value, positions, cos_sin_cache, k_cache, v_cache, slot_mapping)
    lib = load_library()
    T = slot_mapping.shape[0]
    if T == 0:
        return
    if query.dim() != 3 or key.dim() != 3 or value.shape != key.shape or query.stride(2) != 1 or key.stride(2) != 1 or value.stride(2) != 1:
        raise B200AttnError("query/key/value must be [T, H, D] with contiguous head_dim")
    if query.shape[0] < T or key.shape[0] < T or positions.shape[0] < T:
        raise B200AttnError("query/key/value/positions must cover every slot_mapping entry")
    if slot_mapping.dtype != torch.int64 or not slot_mapping.is_contiguous() or positions.dtype != torch.int64 or not positions.is_contiguous():
        raise B200AttnError("slot_mapping and positions must be contiguous int64")
    if cos_sin_cache.dtype != query.dtype or key.dtype != query.dtype or value.dtype != query.dtype or not cos_sin_cache.is_contiguous():
        raise B200AttnError("cos_sin_cache, key and value must have the query dtype (cos_sin_cache contiguous)")
    Hq, Hkv, D = query.shape[1], key.shape[1], key.shape[2]
    sb, ss, sh = _cache_strides(k_cache, v_cache)
    kvdt = _dt(k_cache)
    sstride = _scale_stride(k_scale, Hkv, v_scale) if kvdt == DT_FP8_E4M3 else 0
    check(lib.b200attn_rope_kv_write(
        query.data_ptr(), query.stride(0), query.stride(1), Hq, key.data_ptr(), value.data_ptr(), key.stride(0), key.stride(1),
        value.stride(0), value.stride(1), positions.data_ptr(), cos_sin_cache.data_ptr(), cos_sin_cache.shape[1], (1 if is_neox else 0) | (2 if round_each_op else 0),
        k_cache.data_ptr(), v_cache.data_ptr(), sb, ss, sh, slot_mapping.data_ptr(), T, Hkv, D, k_cache.shape[1],
        _dt(query), kvdt, k_scale.data_ptr(), v_scale.data_ptr(), sstride, _stream(query)), "b200attn_rope_kv_write")


@rope_kv_write.register_fake
def _(query, key, value, positions, cos_sin_cache, is_neox, k_cache, v_cache, slot_mapping, k_scale, v_scale, round_each_op=False) -> None:
    return None


@torch.library.custom_op("b200attn::norm_rope_kv_write", mutates_args=("query", "key", "k_cache", "v_cache"), device_types="cuda")
def norm_rope_kv_write(query: torch.Tensor, key: torch.Tensor, value: torch.Tensor, positions: torch.Tensor,
                       cos_sin_cache: torch.Tensor, is_neox: bool, q_norm_weight: torch.Tensor, k_norm_weight: torch.Tensor,
                       rms_eps: float, mrope_section: list[int], mrope_interleaved: bool, k_cache: torch.Tensor,
                       v_cache: torch.Tensor, slot_mapping: torch.Tensor, k_scale: torch.Tensor, v_scale: torch.Tensor,
                       round_each_op: bool = False) -> None:
    """q/k RMSNorm + rotary embedding (plain or M-RoPE) + KV write in one launch (b200attn_norm_rope_kv_write).
    positions: int64 [T] or [3, T] (then `mrope_section` = [t, h, w] frequency counts, sum = rotary_dim / 2);
    q_norm_weight / k_norm_weight: [D] in the activation dtype, or EMPTY tensors for no normalisation.  Rest as rope_kv_write."""
    _require_cuda(query, key, value, positions, cos_sin_cache, k_cache, v_cache, slot_mapping)
    lib = load_library()
    T = slot_mapping.shape[0]
    if T == 0:
        return
    if query.dim() != 3 or key.dim() != 3 or value.shape != key.shape or query.stride(2) != 1 or key.stride(2) != 1 or value.stride(2) != 1:
        raise B200AttnError("query/key/value must be [T, H, D] with contiguous head_dim")
    if positions.dtype != torch.int64 or positions.dim() not in (1, 2) or positions.stride(-1) != 1 or positions.shape[-1] < T:
        raise B200AttnError("positions must be int64 [T] or [3, T] with contiguous rows")
    if positions.dim() == 2 and (positions.shape[0] != 3 or len(mrope_section) != 3):
        raise B200AttnError("2-D positions must be [3, T] and come with mrope_section = [t, h, w]")
    if query.shape[0] < T or key.shape[0] < T:
        raise B200AttnError("query/key/value must cover every slot_mapping entry")
    if slot_mapping.dtype != torch.int64 or not slot_mapping.is_contiguous():
        raise B200AttnError("slot_mapping must be contiguous int64")
    if cos_sin_cache.dtype != query.dtype or key.dtype != query.dtype or value.dtype != query.dtype or not cos_sin_cache.is_contiguous():
        raise B200AttnError("cos_sin_cache, key and value must have the query dtype (cos_sin_cache contiguous)")
    Hq, Hkv, D = query.shape[1], key.shape[1], key.shape[2]
    qw = q_norm_weight if q_norm_weight.numel() > 0 else None
    kw = k_norm_weight if k_norm_weight.numel() > 0 else None
    for w in (qw, kw):
        if w is not None and (w.dtype != query.dtype or w.numel() != D or not w.is_contiguous() or not w.is_cuda):
            raise B200AttnError("norm weights must be contiguous CUDA tensors of [head_dim] in the query dtype")
    sb, ss, sh = _cache_strides(k_cache, v_cache)
    kvdt = _dt(k_cache)
    sstride = _scale_stride(k_scale, Hkv, v_scale) if kvdt == DT_FP8_E4M3 else 0
    mrope = positions.dim() == 2
    sec = [int(x) for x in mrope_section] if mrope else [0, 0, 0]
    flags = (1 if is_neox else 0) | (2 if round_each_op else 0) | (4 if (mrope and mrope_interleaved) else 0)
    check(lib.b200attn_norm_rope_kv_write(
        query.data_ptr(), query.stride(0), query.stride(1), Hq, key.data_ptr(), value.data_ptr(), key.stride(0), key.stride(1),
        value.stride(0), value.stride(1), positions.data_ptr(), positions.stride(0) if mrope else 0, sec[0], sec[1], sec[2],
        cos_sin_cache.data_ptr(), cos_sin_cache.shape[1], flags, qw.data_ptr() if qw is not None else None,
        kw.data_ptr() if kw is not None else None, float(rms_eps), k_cache.data_ptr(), v_cache.data_ptr(), sb, ss, sh,
        slot_mapping.data_ptr(), T, Hkv, D, k_cache.shape[1], _dt(query), kvdt, k_scale.data_ptr(), v_scale.data_ptr(), sstride,
        _stream(query)), "b200attn_norm_rope_kv_write")


@norm_rope_kv_write.register_fake
def _(query, key, value, positions, cos_sin_cache, is_neox, q_norm_weight, k_norm_weight, rms_eps, mrope_section, mrope_interleaved,
      k_cache, v_cache, slot_mapping, k_scale, v_scale, round_each_op=False) -> None:
    return None


# ------------------------------------------------------------------------------------------------
# decode_plan (once per step) and paged_decode (= split-KV decode kernel + merge kernel)
# ------------------------------------------------------------------------------------------------
@torch.library.custom_op("b200attn::decode_plan", mutates_args=("plan",), device_types="cuda")
def decode_plan(seq_lens: torch.Tensor, plan: torch.Tensor, num_q_heads: int, num_kv_heads: int, num_ctas: int,
                block_table: torch.Tensor, block_size: int, row_map: torch.Tensor) -> None:
    """`block_table` / `block_size` / `row_map` (EMPTY = identity) are the step metadata the decode calls of this step will
    get: the plan copies each CTA's first block ids into its entry (include/b200attn.h)."""
    _require_cuda(seq_lens, plan, block_table)
    if seq_lens.dtype != torch.int32 or plan.dtype != torch.int32 or not seq_lens.is_contiguous():
        raise B200AttnError("seq_lens and plan must be contiguous int32")
    if block_table.dtype != torch.int32 or block_table.dim() != 2 or block_table.stride(1) != 1:
        raise B200AttnError("block_table must be int32 [rows, max_blocks] with contiguous rows")
    rm = row_map if row_map.numel() > 0 else None
    if rm is not None and (rm.dtype != torch.int32 or not rm.is_contiguous() or rm.shape[0] != seq_lens.shape[0]):
        raise B200AttnError("row_map must be contiguous int32 [num_rows]")
    check(load_library().b200attn_decode_plan(seq_lens.data_ptr(), seq_lens.shape[0], num_q_heads, num_kv_heads, num_ctas,
                                              block_table.data_ptr(), block_table.stride(0),
                                              rm.data_ptr() if rm is not None else None, int(block_size),
                                              plan.data_ptr(), plan.numel() * 4, _stream(plan)), "b200attn_decode_plan")


@decode_plan.register_fake
def _(seq_lens, plan, num_q_heads, num_kv_heads, num_ctas, block_table, block_size, row_map) -> None:
    return None


def _decode_args(query, k_cache, v_cache, block_table, seq_lens, out, k_scale, v_scale):
    if query.dim() != 3 or out.dim() != 3 or query.stride(2) != 1 or out.stride(2) != 1 or query.dtype != out.dtype:
        raise B200AttnError("query/out must be [B, Hq, D] of the same dtype with contiguous head_dim")
    if block_table.dtype != torch.int32 or seq_lens.dtype != torch.int32 or block_table.stride(1) != 1:
        raise B200AttnError("block_table / seq_lens must be int32 (block_table rows contiguous)")
    Hkv = k_cache.shape[2]
    kvdt = _dt(k_cache)
    return _cache_strides(k_cache, v_cache), kvdt, (_scale_stride(k_scale, Hkv, v_scale) if kvdt == DT_FP8_E4M3 else 0)


def launch_split_decode(query, k_cache, v_cache, block_table, seq_lens, out, plan, workspace, k_scale, v_scale, sm_scale,
                        num_ctas, lse=None, row_map=None, fresh_tokens=-1, fused_merge=False) -> None:
    """Kernel 2 (b200attn_decode): final rows for unsplit sequences; split ones leave (o, m, l) partials that are merged
    in-kernel by the last-arriving CTA (`fused_merge`) or by a following launch_merge."""
    (sb, ss, sh), kvdt, sstride = _decode_args(query, k_cache, v_cache, block_table, seq_lens, out, k_scale, v_scale)
    if row_map is not None and (row_map.dtype != torch.int32 or not row_map.is_contiguous() or row_map.shape[0] != seq_lens.shape[0]):
        raise B200AttnError("row_map must be contiguous int32 [num_rows]")
    check(load_library().b200attn_decode(
        query.data_ptr(), query.stride(0), query.stride(1), out.data_ptr(), out.stride(0), out.stride(1),
        k_cache.data_ptr(), v_cache.data_ptr(), sb, ss, sh, k_cache.shape[0], block_table.data_ptr(), block_table.stride(0),
        seq_lens.data_ptr(), seq_lens.shape[0], query.shape[1], k_cache.shape[2], query.shape[2], k_cache.shape[1],
        _dt(query), kvdt, float(sm_scale), k_scale.data_ptr(), v_scale.data_ptr(), sstride, plan.data_ptr(), num_ctas,
        workspace.data_ptr(), workspace.numel() * 4, lse.data_ptr() if lse is not None else None,
        row_map.data_ptr() if row_map is not None else None, int(fresh_tokens), DECODE_FUSED_MERGE if fused_merge else 0,
        _stream(query)), "b200attn_decode")


def launch_split_decode_kv_write(query, key, value, slot_mapping, k_cache, v_cache, block_table, seq_lens, out, plan, workspace,
                                 k_scale, v_scale, sm_scale, num_ctas, lse=None, row_map=None, fresh_tokens=1,
                                 fused_merge=False) -> None:
    """b200attn_decode_kv_write: the rows' new tokens are written to the cache by the decode launch itself (FP8 cache, <= 8
    query heads per kv head, rows == tokens) or by a KV-write launch issued from the same call (every other shape)."""
    (sb, ss, sh), kvdt, sstride = _decode_args(query, k_cache, v_cache, block_table, seq_lens, out, k_scale, v_scale)
    n = seq_lens.shape[0]
    if key.dim() != 3 or value.shape != key.shape or key.stride(2) != 1 or value.stride(2) != 1 or key.dtype != query.dtype \
            or value.dtype != query.dtype or key.shape[0] < n:
        raise B200AttnError("key/value must be [rows, Hkv, D] of the query dtype with contiguous head_dim")
    if slot_mapping.dtype != torch.int64 or not slot_mapping.is_contiguous() or slot_mapping.shape[0] < n:
        raise B200AttnError("slot_mapping must be contiguous int64 covering every row")
    if row_map is not None and (row_map.dtype != torch.int32 or not row_map.is_contiguous() or row_map.shape[0] != n):
        raise B200AttnError("row_map must be contiguous int32 [num_rows]")
    check(load_library().b200attn_decode_kv_write(
        query.data_ptr(), query.stride(0), query.stride(1), out.data_ptr(), out.stride(0), out.stride(1),
        key.data_ptr(), value.data_ptr(), key.stride(0), key.stride(1), value.stride(0), value.stride(1), slot_mapping.data_ptr(),
        k_cache.data_ptr(), v_cache.data_ptr(), sb, ss, sh, k_cache.shape[0], block_table.data_ptr(), block_table.stride(0),
        seq_lens.data_ptr(), n, query.shape[1], k_cache.shape[2], query.shape[2], k_cache.shape[1],
        _dt(query), kvdt, float(sm_scale), k_scale.data_ptr(), v_scale.data_ptr(), sstride, plan.data_ptr(), num_ctas,
        workspace.data_ptr(), workspace.numel() * 4, lse.data_ptr() if lse is not None else None,
        row_map.data_ptr() if row_map is not None else None, int(fresh_tokens), DECODE_FUSED_MERGE if fused_merge else 0,
        _stream(query)), "b200attn_decode_kv_write")


def launch_merge(k_cache, seq_lens, out, plan, workspace, v_scale, num_ctas, lse=None) -> None:
    """Kernel 3 only (b200attn_merge); must follow launch_split_decode on the same stream."""
    kvdt = _dt(k_cache)
    Hkv = k_cache.shape[2]
    sstride = _scale_stride(v_scale, Hkv) if kvdt == DT_FP8_E4M3 else 0
    check(load_library().b200attn_merge(
        plan.data_ptr(), num_ctas, workspace.data_ptr(), out.data_ptr(), out.stride(0), out.stride(1), seq_lens.shape[0],
        out.shape[1], Hkv, out.shape[2], _dt(out), v_scale.data_ptr() if kvdt == DT_FP8_E4M3 else None, sstride,
        lse.data_ptr() if lse is not None else None, _stream(out)), "b200attn_merge")


@torch.library.custom_op("b200attn::paged_decode", mutates_args=("out", "plan", "workspace", "lse"), device_types="cuda")
def paged_decode(query: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, block_table: torch.Tensor,
                 seq_lens: torch.Tensor, out: torch.Tensor, plan: torch.Tensor, workspace: torch.Tensor,
                 k_scale: torch.Tensor, v_scale: torch.Tensor, sm_scale: float, num_ctas: int,
                 lse: torch.Tensor, row_map: torch.Tensor, fresh_tokens: int = -1, fused_merge: bool = True) -> None:
    """`lse`: fp32 [B, Hq] to receive the natural-log LSE, or an EMPTY tensor to skip it (a required argument:
    optional mutable tensors trip torch.library's version-counter bump when omitted).  `row_map`: int32 [B] block-table
    row of each decode row, or EMPTY for the identity.  `fresh_tokens`: see include/b200attn.h (-1 = unknown).
    `fused_merge`: merge split-KV partials inside the decode kernel (default) instead of launching the merge kernel.
    `plan` is listed as mutated because the in-kernel merge keeps its arrival counters there."""
    _require_cuda(query, k_cache, v_cache, block_table, seq_lens, out, plan, workspace)
    if seq_lens.shape[0] == 0:
        return
    lse = lse if lse.numel() > 0 else None
    row_map = row_map if row_map.numel() > 0 else None
    # The in-kernel merge folds a pair's partials serially in ONE CTA: fine for the 2-4 partials of a large batch, slow when
    # a few long sequences are cut across the whole grid.  Past ~8 partials per (sequence, kv head) the merge kernel, which
    # spreads them over (pair, head) CTAs, follows instead.  Host-side shape arithmetic only (no device read).
    if query.shape[2] != 128:
        fused_merge = True    # head_dim 64 / 256: the CUDA-core kernel folds its own splits (csrc/decode_simt.cu)
    elif fused_merge and seq_lens.shape[0] * k_cache.shape[2] * 8 < num_ctas:
        fused_merge = False
    launch_split_decode(query, k_cache, v_cache, block_table, seq_lens, out, plan, workspace, k_scale, v_scale, sm_scale,
                        num_ctas, lse, row_map, fresh_tokens, fused_merge)
    if not fused_merge:
        launch_merge(k_cache, seq_lens, out, plan, workspace, v_scale, num_ctas, lse)


@paged_decode.register_fake
def _(query, k_cache, v_cache, block_table, seq_lens, out, plan, workspace, k_scale, v_scale, sm_scale, num_ctas, lse,
      row_map, fresh_tokens=-1, fused_merge=True) -> None:
    return None


@torch.library.custom_op("b200attn::paged_decode_kv_write", mutates_args=("k_cache", "v_cache", "out", "plan", "workspace", "lse"),
                         device_types="cuda")
def paged_decode_kv_write(query: torch.Tensor, key: torch.Tensor, value: torch.Tensor, slot_mapping: torch.Tensor,
                          k_cache: torch.Tensor, v_cache: torch.Tensor, block_table: torch.Tensor, seq_lens: torch.Tensor,
                          out: torch.Tensor, plan: torch.Tensor, workspace: torch.Tensor, k_scale: torch.Tensor,
                          v_scale: torch.Tensor, sm_scale: float, num_ctas: int, lse: torch.Tensor, row_map: torch.Tensor,
                          fresh_tokens: int = 1, fused_merge: bool = True) -> None:
    """KV write + paged decode of a PURE-DECODE step in one call (a backend with forward_includes_kv_cache_update = True):
    key[b] / value[b] go to slot_mapping[b] exactly as torch.ops.b200attn.kv_write stores them, and row b attends to its
    seq_lens[b] positions including that token.  Other arguments as paged_decode."""
    _require_cuda(query, key, value, slot_mapping, k_cache, v_cache, block_table, seq_lens, out, plan, workspace)
    if seq_lens.shape[0] == 0:
        return
    lse = lse if lse.numel() > 0 else None
    row_map = row_map if row_map.numel() > 0 else None
    if query.shape[2] != 128:
        fused_merge = True
    elif fused_merge and seq_lens.shape[0] * k_cache.shape[2] * 8 < num_ctas:
        fused_merge = False
    launch_split_decode_kv_write(query, key, value, slot_mapping, k_cache, v_cache, block_table, seq_lens, out, plan, workspace,
                                 k_scale, v_scale, sm_scale, num_ctas, lse, row_map, fresh_tokens, fused_merge)
    if not fused_merge:
        launch_merge(k_cache, seq_lens, out, plan, workspace, v_scale, num_ctas, lse)


@paged_decode_kv_write.register_fake
def _(query, key, value, slot_mapping, k_cache, v_cache, block_table, seq_lens, out, plan, workspace, k_scale, v_scale, sm_scale,
      num_ctas, lse, row_map, fresh_tokens=1, fused_merge=True) -> None:
    return None


# ------------------------------------------------------------------------------------------------
# paged_prefill: varlen causal attention for rows with q_len > 1 (chunked prefill, mixed batches, spec-decode verification)
# ------------------------------------------------------------------------------------------------
def prefill_tile_tokens(num_q_heads: int, num_kv_heads: int) -> int:
    """Query tokens per 128-row tile of the prefill kernel (128 / query heads per kv head)."""
    n = load_library().b200attn_prefill_tile_tokens(num_q_heads, num_kv_heads)
    if n <= 0:
        raise B200AttnError(f"prefill: unsupported head counts Hq={num_q_heads} Hkv={num_kv_heads}")
    return n


@torch.library.custom_op("b200attn::prefill_plan", mutates_args=("tile_prefix",), device_types="cuda")
def prefill_plan(query_start_loc: torch.Tensor, tile_prefix: torch.Tensor, num_q_heads: int, num_kv_heads: int,
                 min_query_len: int = 1) -> None:
    """tile_prefix[num_reqs + 1] = exclusive scan of each request's query-tile count (0 for requests with fewer than
    `min_query_len` query tokens: the decode kernel's rows in a mixed batch).  Once per scheduler step."""
    _require_cuda(query_start_loc, tile_prefix)
    n = query_start_loc.shape[0] - 1
    if n <= 0:
        return
    if query_start_loc.dtype != torch.int32 or tile_prefix.dtype != torch.int32 or not query_start_loc.is_contiguous() \
            or not tile_prefix.is_contiguous() or tile_prefix.numel() < n + 1:
        raise B200AttnError("query_start_loc / tile_prefix must be contiguous int32 ([num_reqs + 1])")
    check(load_library().b200attn_prefill_plan(query_start_loc.data_ptr(), n, num_q_heads, num_kv_heads, int(min_query_len),
                                               tile_prefix.data_ptr(), _stream(tile_prefix)), "b200attn_prefill_plan")


@prefill_plan.register_fake
def _(query_start_loc, tile_prefix, num_q_heads, num_kv_heads, min_query_len=1) -> None:
    return None


@torch.library.custom_op("b200attn::paged_prefill", mutates_args=("out",), device_types="cuda")
def paged_prefill(query: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, block_table: torch.Tensor,
                  seq_lens: torch.Tensor, query_start_loc: torch.Tensor, tile_prefix: torch.Tensor, out: torch.Tensor,
                  k_scale: torch.Tensor, v_scale: torch.Tensor, sm_scale: float, first_token: int = 0) -> None:
    """query / out [T, Hq, D]; block_table [R, max_blocks], seq_lens [R], query_start_loc [R + 1], tile_prefix [R + 1]
    describe the R requests whose tokens are query[qsl[r] - first_token : qsl[r + 1] - first_token]; token t of a request's
    chunk attends to cache positions [0, seq_len - q_len + t + 1).  FP8 e4m3 cache only."""
    _require_cuda(query, k_cache, v_cache, block_table, seq_lens, query_start_loc, tile_prefix, out)
    R, T = seq_lens.shape[0], query.shape[0]
    if R == 0 or T == 0:
        return
    (sb, ss, sh), kvdt, sstride = _decode_args(query, k_cache, v_cache, block_table, seq_lens, out, k_scale, v_scale)
    if query_start_loc.dtype != torch.int32 or tile_prefix.dtype != torch.int32 or query_start_loc.shape[0] < R + 1 \
            or tile_prefix.numel() < R + 1 or not query_start_loc.is_contiguous() or not seq_lens.is_contiguous():
        raise B200AttnError("query_start_loc / tile_prefix must be contiguous int32 [num_reqs + 1]")
    if out.shape[0] < T or block_table.shape[0] < R:
        raise B200AttnError("out / block_table must cover every query token / request")
    check(load_library().b200attn_prefill(
        query.data_ptr(), query.stride(0), query.stride(1), out.data_ptr(), out.stride(0), out.stride(1),
        k_cache.data_ptr(), v_cache.data_ptr(), sb, ss, sh, k_cache.shape[0], block_table.data_ptr(), block_table.stride(0),
        seq_lens.data_ptr(), query_start_loc.data_ptr(), R, T, int(first_token), query.shape[1], k_cache.shape[2], query.shape[2],
        k_cache.shape[1], _dt(query), kvdt, float(sm_scale), k_scale.data_ptr(), v_scale.data_ptr(), sstride,
        tile_prefix.data_ptr(), _stream(query)), "b200attn_prefill")


@paged_prefill.register_fake
def _(query, k_cache, v_cache, block_table, seq_lens, query_start_loc, tile_prefix, out, k_scale, v_scale, sm_scale,
      first_token=0) -> None:
    return None


# ------------------------------------------------------------------------------------------------
# merge_lse: combine P attention results by log-sum-exp (sequence-split across shards / ranks)
# ------------------------------------------------------------------------------------------------
@torch.library.custom_op("b200attn::merge_lse", mutates_args=("out", "lse_out"), device_types="cuda")
def merge_lse(parts_o: torch.Tensor, parts_lse: torch.Tensor, out: torch.Tensor, lse_out: torch.Tensor) -> None:
    """parts_o [P, n, Hq, D] (bf16/fp16), parts_lse [P, n, Hq] fp32 natural log; out [n, Hq, D]; lse_out [n, Hq] or empty."""
    _require_cuda(parts_o, parts_lse, out)
    lse_out = lse_out if lse_out.numel() > 0 else None
    if not (parts_o.is_contiguous() and parts_lse.is_contiguous() and out.is_contiguous()):
        raise B200AttnError("merge_lse needs contiguous tensors")
    P, n, Hq, D = parts_o.shape
    check(load_library().b200attn_merge_lse(
        parts_o.data_ptr(), parts_lse.data_ptr(), P, parts_o.stride(0), parts_lse.stride(0), out.data_ptr(),
        lse_out.data_ptr() if lse_out is not None else None, n, Hq, D, _dt(out), _stream(out)), "b200attn_merge_lse")


@merge_lse.register_fake
def _(parts_o, parts_lse, out, lse_out) -> None:
    return None


def aligned_grid(max_ctas: int, num_seqs: int, num_q_heads: int, num_kv_heads: int) -> int:
    """Decode grid for a batch (host-side shape arithmetic, capture-safe).  When the (sequence, kv head) pairs are few
    enough that each can get k >= 2 CTAs, the grid is rounded down to pairs * k (if that keeps >= 80 % of the resident
    CTAs): with equal-length sequences every CTA then owns exactly 1/k of ONE pair -- a single segment, a k-way merge and
    no mid-range segment switch (c2: 296 -> 256 CTAs, +2.5 %; 32 x 4k: +14 %).  Ragged batches are still cut into equal
    tile counts by the plan kernel, whatever the grid."""
    group = max(1, num_q_heads // max(1, num_kv_heads))
    pairs = num_seqs * num_kv_heads * (-(-group // 16))
    if pairs > 0:
        k = max_ctas // pairs
        # an even k keeps every cluster of 2 inside one pair, so half of each pair's merge goes through distributed shared
        # memory (thinker at tp 4, 32 pairs: 9 -> 8 CTAs per pair, 49.8 -> 48.9 us)
        if k > 2 and k % 2 == 1 and pairs * (k - 1) * 10 >= max_ctas * 8:
            k -= 1
        if k >= 2 and pairs * k * 10 >= max_ctas * 8:
            return pairs * k
    return max_ctas


# ------------------------------------------------------------------------------------------------
# Convenience wrapper owning the plan / workspace buffers (what the metadata builder keeps per device)
# ------------------------------------------------------------------------------------------------
class DecodeWorkspace:
    """Pre-allocated, capture-safe buffers for one device: plan (int32) + split-KV partials (fp32)."""

    def __init__(self, device, max_num_seqs: int, kv_dtype: torch.dtype, head_dim: int = 128, num_ctas: int | None = None,
                 group_size: int = 8):
        """`group_size` = query heads per kv head: 16-bit KV with more than 8 (the two-accumulator-set variant of the
        kernel, one resident CTA per SM) gets half the default grid so the persistent CTAs still run as one wave."""
        self.device = torch.device(device)
        with torch.cuda.device(self.device):
            self.num_ctas = int(num_ctas) if num_ctas else default_ctas(kv_dtype)
        if not num_ctas and kv_dtype not in _FP8_DTYPES and group_size > 8:
            self.num_ctas = max(1, self.num_ctas // 2)
        self._fixed_grid = bool(num_ctas)   # an explicit grid (tests, sweeps) is used as given
        self.grid = self.num_ctas           # grid of the current plan (<= num_ctas, which sizes the workspace)
        self.max_num_seqs = max_num_seqs
        self.plan = torch.zeros(plan_numel(max_num_seqs), dtype=torch.int32, device=self.device)
        self.workspace = torch.empty(workspace_numel(head_dim, self.num_ctas), dtype=torch.float32, device=self.device)
        self._no_lse = torch.empty(0, dtype=torch.float32, device=self.device)
        self._no_map = torch.empty(0, dtype=torch.int32, device=self.device)

    def grid_for(self, num_seqs: int, num_q_heads: int, num_kv_heads: int) -> int:
        """Decode grid of the next plan: `aligned_grid` unless the workspace was built with an explicit grid."""
        if self._fixed_grid:
            return self.num_ctas
        return aligned_grid(self.num_ctas, num_seqs, num_q_heads, num_kv_heads)

    def make_plan(self, seq_lens: torch.Tensor, num_q_heads: int, num_kv_heads: int, block_table: torch.Tensor,
                  block_size: int, row_map: torch.Tensor | None = None) -> None:
        """Once per step (shared by all layers).  `block_table` / `row_map` must be the tensors `decode` gets afterwards."""
        if seq_lens.shape[0] > self.max_num_seqs:
            raise B200AttnError(f"batch {seq_lens.shape[0]} exceeds planned max_num_seqs {self.max_num_seqs}")
        self.grid = self.grid_for(seq_lens.shape[0], num_q_heads, num_kv_heads)
        torch.ops.b200attn.decode_plan(seq_lens, self.plan, num_q_heads, num_kv_heads, self.grid, block_table, block_size,
                                       self._no_map if row_map is None else row_map)

    def decode_kv_write(self, query, key, value, slot_mapping, k_cache, v_cache, block_table, seq_lens, out, k_scale, v_scale,
                        sm_scale, lse=None, row_map=None, fresh_tokens: int = 1, fused_merge: bool = True) -> None:
        torch.ops.b200attn.paged_decode_kv_write(query, key, value, slot_mapping, k_cache, v_cache, block_table, seq_lens, out,
                                                 self.plan, self.workspace, k_scale, v_scale, sm_scale, self.grid,
                                                 self._no_lse if lse is None else lse,
                                                 self._no_map if row_map is None else row_map, fresh_tokens, fused_merge)

    def decode_per_token_head(self, query, k_cache, v_cache, k_scale_cache, v_scale_cache, block_table, seq_lens, out, sm_scale,
                              lse=None, row_map=None, fresh_tokens: int = -1) -> None:
        torch.ops.b200attn.paged_decode_per_token_head(query, k_cache, v_cache, k_scale_cache, v_scale_cache, block_table, seq_lens,
                                                       out, self.plan, self.workspace, sm_scale, self.grid,
                                                       self._no_lse if lse is None else lse,
                                                       self._no_map if row_map is None else row_map, fresh_tokens)

    def decode(self, query, k_cache, v_cache, block_table, seq_lens, out, k_scale, v_scale, sm_scale, lse=None,
               row_map=None, fresh_tokens: int = -1, fused_merge: bool = True) -> None:
        torch.ops.b200attn.paged_decode(query, k_cache, v_cache, block_table, seq_lens, out, self.plan, self.workspace,
                                        k_scale, v_scale, sm_scale, self.grid, self._no_lse if lse is None else lse,
                                        self._no_map if row_map is None else row_map, fresh_tokens, fused_merge)
