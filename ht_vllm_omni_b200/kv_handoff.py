"""KV hand-off gather: one request's blocks out of every layer's paged cache, one kernel launch and
one pinned device->host copy instead of the reference's per-layer index / flatten / .cpu() loop.

Mirrors `OmniKVTransferManager._extract_kv_cache` ($REF/vllm_omni/distributed/omni_connectors/kv_transfer_manager.py:224-301)
and `normalize_layer_kv` ($REF/.../utils/kv_utils.py:13-85): same arguments, same result layout
(`key_cache[l]`, `value_cache[l]` = CPU tensors `[min(seq_len, n_valid * block_size), n_heads, head_dim]` of the cache
dtype, raw fp8 bytes included), same filtering of out-of-range block ids, `None` when no layer yields data.
There is no CPU fallback: the caches must live on a CUDA device and libb200attn.so must be built.
"""
from __future__ import annotations

from typing import Any, Sequence

import torch

from ._lib import B200AttnError, check, load_library


def normalize_layer_kv(layer_kv) -> tuple[torch.Tensor, torch.Tensor] | None:
    """(key_blocks, value_blocks) views `[num_blocks, block_size, n_heads, head_dim]` of one layer's cache, for the stacked
    `[2, nb, ...]` / `[nb, 2, ...]` tensors and the (key, value) tuple the reference accepts (kv_utils.py:54-85)."""
    if isinstance(layer_kv, torch.Tensor):
        if layer_kv.ndim >= 3 and layer_kv.shape[0] == 2:
            return layer_kv[0], layer_kv[1]
        if layer_kv.ndim >= 3 and layer_kv.shape[1] == 2:
            return layer_kv[:, 0], layer_kv[:, 1]
        return None
    if isinstance(layer_kv, tuple) and len(layer_kv) == 2 and all(isinstance(t, torch.Tensor) for t in layer_kv):
        k, v = layer_kv
        return (k, v) if k.ndim >= 2 and v.ndim >= 2 else None
    return None


def gather_request_kv(block_ids: Sequence[int], seq_len: int, kv_caches: Sequence[Any]) -> tuple[list, list] | None:
    """The gather itself: returns (key_cache, value_cache), lists over layers of CPU tensors (views of ONE pinned staging
    buffer) or None entries for layers without a usable cache; None when nothing could be extracted."""
    num_layers = len(kv_caches)
    pairs = [normalize_layer_kv(kv) for kv in kv_caches]
    live = [i for i, p in enumerate(pairs) if p is not None and p[0].ndim == 4]
    if not live:
        return None
    k0 = pairs[live[0]][0]
    if not k0.is_cuda:
        raise B200AttnError("kv hand-off gather: the KV caches must be CUDA tensors (no CPU fallback)")
    nb, bs, H, D = k0.shape
    es = k0.element_size()
    strides = (k0.stride(0), k0.stride(1), k0.stride(2))
    for i in live:
        k, v = pairs[i]
        if k.shape != k0.shape or v.shape != k0.shape or k.dtype != k0.dtype or v.dtype != k0.dtype or k.stride(3) != 1 or \
                v.stride(3) != 1 or (k.stride(0), k.stride(1), k.stride(2)) != strides or (v.stride(0), v.stride(1), v.stride(2)) != strides:
            raise B200AttnError("kv hand-off gather: every layer's K and V plane must share shape, dtype and strides")
    # the reference keeps only ids inside [0, num_blocks) (kv_transfer_manager.py:267-270) and truncates to seq_len (:280-282)
    valid = [int(b) for b in block_ids if 0 <= int(b) <= nb - 1]
    if not valid:
        return None
    n_tokens = max(0, min(int(seq_len), len(valid) * bs))
    dev = k0.device
    key_cache: list = [None] * num_layers
    value_cache: list = [None] * num_layers
    L = len(live)
    # straight from torch's caching pinned-host allocator (no pageable detour; blocks of finished requests are reused)
    host = torch.empty((L, 2, n_tokens, H, D), dtype=k0.dtype, pin_memory=True)
    if n_tokens > 0:
        with torch.cuda.device(dev):
            ids = torch.tensor(valid, dtype=torch.int32).pin_memory().to(dev, non_blocking=True)
            kp = torch.tensor([pairs[i][0].data_ptr() for i in live], dtype=torch.int64).pin_memory().to(dev, non_blocking=True)
            vp = torch.tensor([pairs[i][1].data_ptr() for i in live], dtype=torch.int64).pin_memory().to(dev, non_blocking=True)
            out = torch.empty((L, 2, n_tokens, H, D), dtype=k0.dtype, device=dev)
            stream = torch.cuda.current_stream(dev)
            check(load_library().b200attn_kv_gather(kp.data_ptr(), vp.data_ptr(), L, strides[0] * es, strides[1] * es, strides[2] * es,
                                                    nb, ids.data_ptr(), n_tokens, H, bs, D * es, out.data_ptr(), stream.cuda_stream),
                  "b200attn_kv_gather")
            host.copy_(out, non_blocking=True)
            stream.synchronize()
    for j, i in enumerate(live):
        key_cache[i], value_cache[i] = host[j, 0], host[j, 1]
    return key_cache, value_cache


def kv_scale_metadata(kv_scales: Sequence[Any]) -> dict[str, list]:
    """FP8 caches travel as raw e4m3 bytes; the receiving stage needs the scales to read them (value = byte * scale[head]).
    `kv_scales`: per layer a (k_scale, v_scale) pair -- the attention layer's `_k_scale` / `_v_scale` tensors (1 element =
    per-tensor, Hkv elements = per-head) or floats -- or None for a layer without a cache.  Returns plain lists of floats per
    layer (JSON / msgpack friendly, like the rest of the metadata), fetched with ONE device->host copy."""
    flat, shape = [], []
    for pair in kv_scales:
        if pair is None:
            shape.append(None)
            continue
        k, v = (torch.as_tensor(x, dtype=torch.float32).reshape(-1) for x in pair)
        shape.append((k.numel(), v.numel()))
        flat += [k, v]
    if not flat:
        return {"k_scale": [None] * len(shape), "v_scale": [None] * len(shape)}
    dev = next((t.device for t in flat if t.is_cuda), flat[0].device)
    host = torch.cat([t.to(dev) for t in flat]).cpu().tolist()
    ks, vs, o = [], [], 0
    for sh in shape:
        if sh is None:
            ks.append(None); vs.append(None)
            continue
        ks.append(host[o:o + sh[0]]); o += sh[0]
        vs.append(host[o:o + sh[1]]); o += sh[1]
    return {"k_scale": ks, "v_scale": vs}


def extract_kv_cache(req_id: str, block_ids: list[int], seq_len: int, kv_caches: list, block_size: int, cache_dtype: str,
                     custom_metadata: dict[str, Any] | None = None, kv_scales: Sequence[Any] | None = None) -> dict | None:
    """Drop-in for `_extract_kv_cache`: the reference wraps the same fields in its KVCacheTransferData dataclass
    (request_id, layer_blocks, block_ids, metadata); INTEGRATION.md shows the two-line change that makes it call this.
    `kv_scales` (one (k_scale, v_scale) per layer, see `kv_scale_metadata`) adds `metadata["kv_scales"]` so an FP8 cache's raw
    bytes can be dequantised downstream ("raw bytes + scales in metadata"); the reference's own fields and
    their meaning are unchanged, and callers that pass nothing get exactly the reference's metadata."""
    got = gather_request_kv(block_ids, seq_len, kv_caches)
    if got is None:
        return None
    key_cache, value_cache = got
    meta = {"block_size": block_size, "num_layers": len(kv_caches), "dtype": str(cache_dtype), "seq_len": seq_len}
    if kv_scales is not None:
        if len(kv_scales) != len(kv_caches):
            raise B200AttnError(f"kv_scales must have one entry per layer ({len(kv_scales)} vs {len(kv_caches)})")
        meta["kv_scales"] = kv_scale_metadata(kv_scales)
    meta.update(custom_metadata or {})
    return {"request_id": req_id, "layer_blocks": {"key_cache": key_cache, "value_cache": value_cache}, "block_ids": block_ids,
            "metadata": meta}
