"""CPU ORACLE (test infrastructure only) for the paged KV write and paged decode attention hot path.

    *** This file is a checker.  Nothing under ht_vllm_omni_b200/ may import or call it; only tests/,
    *** __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs do.

It is a numpy restatement of the algorithm the reference (ht-vllm-omni) runs for this path.  The reference has no
arithmetic of its own here: it delegates to the un-vendored dependency **vLLM** (pinned ``vllm==0.18.0`` in
docs/getting_started/quickstart.md:22; this image carries vllm 0.22.0), so each function cites the vLLM file
(``$VLLM`` = site-packages/vllm) whose published algorithm it follows, and the reference call site that reaches it:

  kv_write_ref          $VLLM/v1/attention/ops/triton_reshape_and_cache_flash.py:44-124
                        (called via vllm_omni/worker/gpu_ar_model_runner.py:228-298 -> unified_kv_cache_update)
  paged_decode_ref      $VLLM/v1/attention/ops/triton_unified_attention.py:38-55 (dequant), :395-545 (tile loop),
                        triton_attention_helpers.py:361-383 (softmax_step), :606-643 (2D epilogue)
  paged_decode_split_ref  same kernel, IS_3D path :297-300,:547-600 + reduce_segments :645-733
  paged_prefill_ref     same kernel with q_len > 1: :286-303 (causal length per query row), :511-520 (mask)
  merge_lse_ref         $VLLM/v1/attention/ops/merge_attn_states.py:7-101

Pinning: the reference's own tests hold NO golden vector for this boundary.  The oracle is pinned
instead against outputs of vLLM 0.22.0's own Triton kernels executed in this container with TRITON_INTERPRET=1
(tests/golden/make_golden.py -> tests/golden/*.npz; fp16 activations, because the Triton interpreter mishandles
bf16), and on the GPU box the CUDA path is additionally compared with vLLM's compiled / Triton / FlashInfer
kernels directly.  Saturation and NaN behaviour of the fp8 cast come from the CUDA op
(`__nv_cvt_float_to_fp8(x, __NV_SATFINITE, __NV_E4M3)`) and are only pinned on the GPU box.
"""
from __future__ import annotations

import numpy as np

# ------------------------------------------------------------------------------------------------
# number formats
# ------------------------------------------------------------------------------------------------


def _e4m3_table() -> np.ndarray:
    t = np.zeros(256, dtype=np.float32)
    for b in range(256):
        s = -1.0 if b & 0x80 else 1.0
        e = (b >> 3) & 0xF
        m = b & 7
        if e == 0xF and m == 7:
            v = np.nan
        elif e == 0:
            v = (m / 8.0) * 2.0 ** -6
        else:
            v = (1.0 + m / 8.0) * 2.0 ** (e - 7)
        t[b] = s * v
    return t


E4M3_LUT = _e4m3_table()           # byte -> float32 (0x7f / 0xff are NaN)
_POS = E4M3_LUT[:127].astype(np.float64)        # 0 .. 448 ascending, codes 0..126
_MID = (_POS[:-1] + _POS[1:]) / 2.0             # rounding boundaries


def e4m3_decode(b: np.ndarray) -> np.ndarray:
    return E4M3_LUT[np.asarray(b, dtype=np.uint8)]


def e4m3_encode_satfinite(x: np.ndarray) -> np.ndarray:
    """float32 -> e4m3fn byte: round-to-nearest-even, saturate to +-448, NaN -> 0x7f | sign."""
    x = np.asarray(x, dtype=np.float32)
    a = np.abs(x).astype(np.float64)
    sign = (np.signbit(x)).astype(np.uint8) << 7
    idx = np.searchsorted(_MID, a, side="left")          # a <= MID[idx] -> candidate code idx (ties land here)
    idx = np.minimum(idx, 126)
    tie = (idx < 126) & (a == _MID[np.minimum(idx, 125)])
    idx = np.where(tie & (idx % 2 == 1), idx + 1, idx)   # ties to even mantissa
    idx = np.minimum(idx, 126).astype(np.uint8)
    out = idx | sign
    return np.where(np.isnan(x), np.uint8(0x7F) | sign, out).astype(np.uint8)


def bf16_round(x: np.ndarray) -> np.ndarray:
    """float32 value rounded to the nearest bf16 (ties to even), returned as float32."""
    u = np.asarray(x, dtype=np.float32).view(np.uint32)
    r = ((u >> 16) & 1) + np.uint32(0x7FFF)
    y = ((u + r) & np.uint32(0xFFFF0000)).view(np.float32)
    return np.where(np.isnan(x), x, y).astype(np.float32)


def act_round(x: np.ndarray, act: str) -> np.ndarray:
    if act == "bf16":
        return bf16_round(x)
    if act == "fp16":
        return np.asarray(x, dtype=np.float32).astype(np.float16).astype(np.float32)
    if act == "fp32":
        return np.asarray(x, dtype=np.float32)
    raise ValueError(act)


def bf16_bits_to_f32(u16: np.ndarray) -> np.ndarray:
    return (np.asarray(u16, dtype=np.uint16).astype(np.uint32) << 16).view(np.float32)


def f32_to_bf16_bits(x: np.ndarray) -> np.ndarray:
    return (bf16_round(x).view(np.uint32) >> 16).astype(np.uint16)


# ------------------------------------------------------------------------------------------------
# KV write  (triton_reshape_and_cache_flash.py:44-124)
# ------------------------------------------------------------------------------------------------


def kv_write_ref(key, value, k_cache, v_cache, slot_mapping, k_scale=None, v_scale=None, fp8=False):
    """In place.  key/value float32 arrays [T, Hkv, D] holding exactly-representable activation values;
    k_cache/v_cache [nb, bs, Hkv, D]: uint8 (fp8 bytes) when fp8 else float32 holding 16-bit values.
    Scales: scalar or [Hkv] float32 (per-head extension).  slot < 0 -> skipped (:46-48)."""
    bs = k_cache.shape[1]
    Hkv = key.shape[1]
    ks = np.broadcast_to(np.asarray(1.0 if k_scale is None else k_scale, dtype=np.float32).reshape(-1), (Hkv,)) if fp8 else None
    vs = np.broadcast_to(np.asarray(1.0 if v_scale is None else v_scale, dtype=np.float32).reshape(-1), (Hkv,)) if fp8 else None
    for t, slot in enumerate(np.asarray(slot_mapping).tolist()):
        if slot < 0:
            continue
        blk, off = slot // bs, slot % bs                      # :50-51
        if fp8:
            # x / scale in float32 (:93 ``key_load / tl.load(k_scale)``), then the implicit fp8 store cast (:113-118)
            k_cache[blk, off] = e4m3_encode_satfinite(key[t].astype(np.float32) / ks[:, None])
            v_cache[blk, off] = e4m3_encode_satfinite(value[t].astype(np.float32) / vs[:, None])
        else:
            k_cache[blk, off] = key[t]
            v_cache[blk, off] = value[t]


def kv_write_per_token_head_ref(key, value, k_cache, v_cache, k_scale_cache, v_scale_cache, slot_mapping):
    """_reshape_cache_per_token_head ($VLLM/v1/attention/ops/triton_reshape_and_cache_flash.py:138-236), FP8 flavour
    (QUANT_MAX 448): per (token, head)  scale = max(max|x| / 448, 1e-6) (:188, :214), stored in the float32 side cache
    [nb, bs, Hkv] (:189-195), bytes = e4m3(clamp(x * (1 / scale), -448, 448)) (:197-206).  IEEE float32 arithmetic -- what
    numpy and the Triton interpreter compute; the compiled Triton kernel's `/` is div.full.f32 (<= 2 ulp), which the CUDA
    kernel reproduces by default and the on-box test pins against the compiled reference.  In place."""
    bs = k_cache.shape[1]
    for t, slot in enumerate(np.asarray(slot_mapping).tolist()):
        if slot < 0:
            continue
        blk, off = slot // bs, slot % bs
        for src, cache, sc in ((key, k_cache, k_scale_cache), (value, v_cache, v_scale_cache)):
            x = np.asarray(src[t], dtype=np.float32)                                  # [Hkv, D]
            scale = np.maximum(np.abs(x).max(axis=-1) / np.float32(448.0), np.float32(1e-6)).astype(np.float32)
            sc[blk, off] = scale
            inv = (np.float32(1.0) / scale).astype(np.float32)
            cache[blk, off] = e4m3_encode_satfinite(np.clip((x * inv[:, None]).astype(np.float32), -448.0, 448.0))


def dequant_per_token_head(cache_u8, scale_cache) -> np.ndarray:
    """float32 [nb, bs, Hkv, D] = e4m3(byte) * scale[block, slot, head]: what kernel_unified_attention sees in
    FP8_PER_TOKEN_HEAD mode (triton_unified_attention.py:472-545 applies k_scale[token] to S and v_scale[token] to P, which
    is this product by linearity).  Feed it to paged_decode_exact(fp8=False) for the attention yardstick."""
    return (e4m3_decode(cache_u8).astype(np.float32) * np.asarray(scale_cache, dtype=np.float32)[..., None]).astype(np.float32)


# ------------------------------------------------------------------------------------------------
# RoPE + KV write  (AttentionImpl.do_rope_and_kv_cache_update, $VLLM/v1/attention/backend.py:815-840)
# Unfused reference = RotaryEmbedding.forward ($VLLM/model_executor/layers/rotary_embedding/base.py:130-180,
# common.py:144-183 ApplyRotaryEmb.forward_static; csrc/pos_encoding_kernels.cu computes the same scalar_t arithmetic)
# followed by reshape_and_cache_flash.
# ------------------------------------------------------------------------------------------------


def rope_cos_sin_cache(rotary_dim: int, max_pos: int, base: float, act: str) -> np.ndarray:
    """[max_pos, rotary_dim] = cat(cos, sin) of outer(t, base^(-2i/rotary_dim)), rounded to the activation dtype
    (base.py:_compute_inv_freq / _compute_cos_sin_cache + _match_cos_sin_cache_dtype)."""
    inv_freq = (1.0 / (np.float32(base) ** (np.arange(0, rotary_dim, 2, dtype=np.float32) / np.float32(rotary_dim)))).astype(np.float32)
    freqs = np.outer(np.arange(max_pos, dtype=np.float32), inv_freq).astype(np.float32)
    return act_round(np.concatenate([np.cos(freqs), np.sin(freqs)], axis=-1).astype(np.float32), act)


def rope_ref(x, positions, cos_sin_cache, rotary_dim: int, is_neox: bool, act: str, each_op: bool = False) -> np.ndarray:
    """x float32 [T, H, D] holding activation-dtype values; returns the rotated copy.
    each_op=False: the compiled CUDA op's arithmetic (csrc/pos_encoding_kernels.cu) -- fp32, one rounding; pinned on the box
    against torch.ops._C.rotary_embedding (tests/test_gpu_vs_vllm.py, tests/tools/rope_probe.py).
    each_op=True: every elementary operation rounded to the activation dtype, as torch does for bf16/fp16 tensors
    (forward_native, common.py:173-174; pinned by tests/golden/vllm_rope_native.npz):
        o1 = rnd(rnd(x1 * cos) - rnd(x2 * sin));  o2 = rnd(rnd(x2 * cos) + rnd(x1 * sin))
    neox: (x1, x2) = halves of the rotary span; GPT-J: even / odd elements (common.py:167-171)."""
    x = np.asarray(x, dtype=np.float32)
    out = x.copy()
    half = rotary_dim // 2
    cs = np.asarray(cos_sin_cache, dtype=np.float32)[np.asarray(positions, dtype=np.int64)]   # [T, rotary_dim]
    cos, sin = cs[:, None, :half], cs[:, None, half:]
    rot = x[..., :rotary_dim]
    if is_neox:
        x1, x2 = rot[..., :half], rot[..., half:]
    else:
        x1, x2 = rot[..., 0::2], rot[..., 1::2]
    r = lambda a: act_round(a, act)  # noqa: E731
    if each_op:
        o1 = r(r(x1 * cos) - r(x2 * sin))
        o2 = r(r(x2 * cos) + r(x1 * sin))
    else:   # products of two 16-bit floats are exact in float32
        o1 = r((x1 * cos).astype(np.float32) - (x2 * sin).astype(np.float32))
        o2 = r((x2 * cos).astype(np.float32) + (x1 * sin).astype(np.float32))
    if is_neox:
        out[..., :half], out[..., half:rotary_dim] = o1, o2
    else:
        out[..., 0:rotary_dim:2], out[..., 1:rotary_dim:2] = o1, o2
    return out


def mrope_axes(rotary_dim: int, mrope_section, interleaved: bool) -> np.ndarray:
    """Position axis (0 temporal, 1 height, 2 width) of every rotary frequency 0 .. rotary_dim/2 - 1
    ($VLLM/model_executor/layers/rotary_embedding/mrope.py: chunked split of forward_native :291-298; interleaved layout
    of apply_interleaved_rope :190-198: x_t[1 : 3h : 3] <- H, x_t[2 : 3w : 3] <- W)."""
    half = rotary_dim // 2
    t, h, w = (int(x) for x in mrope_section)
    assert t + h + w == half
    ax = np.zeros(half, dtype=np.int64)
    if interleaved:
        ax[1:3 * h:3] = 1
        ax[2:3 * w:3] = 2
    else:
        ax[t:t + h] = 1
        ax[t + h:] = 2
    return ax


def mrope_ref(x, positions3, cos_sin_cache, rotary_dim: int, is_neox: bool, act: str, mrope_section, interleaved: bool,
              each_op: bool = False) -> np.ndarray:
    """M-RoPE (MRotaryEmbedding.forward_native, mrope.py:263-322): positions3 int64 [3, T]; frequency f of token t takes its
    cos / sin from cos_sin_cache[positions3[axis(f), t]]; the rotation itself is rope_ref's."""
    x = np.asarray(x, dtype=np.float32)
    half = rotary_dim // 2
    cache = np.asarray(cos_sin_cache, dtype=np.float32)
    p3 = np.asarray(positions3, dtype=np.int64)
    ax = mrope_axes(rotary_dim, mrope_section, interleaved)
    T = p3.shape[1]
    rows = p3[ax[None, :].repeat(T, 0), np.arange(T)[:, None]]            # [T, half]: cache row per (token, frequency)
    f = np.arange(half)[None, :]
    table = np.concatenate([cache[rows, f], cache[rows, half + f]], axis=1)   # per-token [cos | sin] rows
    out = np.empty_like(x)
    for t in range(T):   # rope_ref with a one-row "cache" per token
        out[t:t + 1] = rope_ref(x[t:t + 1], np.zeros(1, np.int64), table[t:t + 1], rotary_dim, is_neox, act, each_op)
    return out


def rms_norm_ref(x, weight, eps: float, act: str) -> np.ndarray:
    """RMSNorm over the last axis as the reference computes it for 16-bit activations (RMSNorm.forward_native,
    $VLLM/model_executor/layers/layernorm.py; csrc/layernorm_kernels.cu rms_norm_kernel): statistics in fp32,
        y = rnd(rnd(x * rsqrt(mean(x^2) + eps)) * w)
    x float32 [..., D] holding activation-dtype values, weight [D] likewise.  The fp32 sum's association order is the
    reference's implementation detail (block reduce / pairwise), so comparisons allow one unit in the last place."""
    x = np.asarray(x, dtype=np.float32)
    var = np.mean(x.astype(np.float64) ** 2, axis=-1, keepdims=True).astype(np.float32)
    inv = (1.0 / np.sqrt(var.astype(np.float64) + np.float64(np.float32(eps)))).astype(np.float32)
    y = act_round((x * inv).astype(np.float32), act)
    return act_round((y * np.asarray(weight, dtype=np.float32)).astype(np.float32), act)


def rope_kv_write_ref(query, key, value, positions, cos_sin_cache, rotary_dim, is_neox, act, k_cache, v_cache, slot_mapping,
                      k_scale=None, v_scale=None, fp8=False, each_op=False):
    """Returns (q', k'); k' and value are written to the cache exactly as kv_write_ref does."""
    q2 = rope_ref(query, positions, cos_sin_cache, rotary_dim, is_neox, act, each_op)
    k2 = rope_ref(key, positions, cos_sin_cache, rotary_dim, is_neox, act, each_op)
    kv_write_ref(k2, value, k_cache, v_cache, slot_mapping, k_scale, v_scale, fp8)
    return q2, k2


# ------------------------------------------------------------------------------------------------
# KV hand-off gather  
# ------------------------------------------------------------------------------------------------


def kv_extract_ref(block_ids, seq_len, key_blocks, value_blocks):
    """One layer of OmniKVTransferManager._extract_kv_cache
    ($REF/vllm_omni/distributed/omni_connectors/kv_transfer_manager.py:266-282): key_blocks / value_blocks are
    [num_blocks, block_size, n_heads, head_dim] arrays of any dtype (raw fp8 bytes included).  Ids outside [0, num_blocks)
    are dropped (:269-270), the selected blocks are flattened over (block, slot) (:276-279) and cut to seq_len (:280-282).
    Returns (flat_k, flat_v) or None when no id is valid (:271-272)."""
    max_block = min(key_blocks.shape[0], value_blocks.shape[0]) - 1
    valid = [b for b in block_ids if 0 <= b <= max_block]
    if not valid:
        return None
    fk = key_blocks[valid].reshape((-1,) + key_blocks.shape[2:])
    fv = value_blocks[valid].reshape((-1,) + value_blocks.shape[2:])
    if seq_len < fk.shape[0]:
        fk, fv = fk[:seq_len], fv[:seq_len]
    return np.ascontiguousarray(fk), np.ascontiguousarray(fv)


# ------------------------------------------------------------------------------------------------
# paged decode attention (q_len = 1)
# ------------------------------------------------------------------------------------------------


def _gather(cache, block_table_row, L, h, bs):
    idx = np.arange(L)
    return cache[block_table_row[idx // bs], idx % bs, h]      # [L, D]


def _dequant(raw, scale, fp8, act):
    if fp8:   # _cast_kv_tile KV_QUANT_MODE == 1: (fp8 -> fp32) * scale -> Q dtype   (:52-54)
        return act_round(e4m3_decode(raw) * np.float32(scale), act)
    return np.asarray(raw, dtype=np.float32)


def _attend_range(q, K, V, sm_scale, act, tile, lo, hi):
    """Online softmax over tokens [lo, hi) in `tile`-sized steps; returns (acc, M, L) un-normalised.
    q [G, D] fp32; K, V [L, D] fp32 (already in activation precision)."""
    G = q.shape[0]
    M = np.full(G, -np.inf, dtype=np.float32)
    Lsum = np.ones(G, dtype=np.float32)                        # L initialised to 1.0 (:351)
    acc = np.zeros((G, V.shape[1]), dtype=np.float32)
    for j in range(lo, hi, tile):
        e = min(j + tile, hi)
        S = np.float32(sm_scale) * (q @ K[j:e].T).astype(np.float32)        # :503-508
        m_j = np.maximum(M, S.max(axis=1))                                  # softmax_step
        m_j = np.where(m_j > -np.inf, m_j, np.float32(0.0))
        P = np.exp(S - m_j[:, None]).astype(np.float32)
        alpha = np.exp(M - m_j).astype(np.float32)
        Lsum = Lsum * alpha + P.sum(axis=1, dtype=np.float32)
        M = m_j
        acc = acc * alpha[:, None] + (act_round(P, act) @ V[j:e]).astype(np.float32)   # P.to(V.dtype) @ V  (:545)
    return acc, M, Lsum


def paged_decode_ref(q, k_cache, v_cache, block_table, seq_lens, sm_scale, k_scale=None, v_scale=None, fp8=False,
                     act="bf16", tile=16, return_lse=False):
    """q float32 [B, Hq, D] (activation-representable).  Returns float32 [B, Hq, D] rounded to `act`."""
    B, Hq, D = q.shape
    bs, Hkv = k_cache.shape[1], k_cache.shape[2]
    G = Hq // Hkv
    ks = np.broadcast_to(np.asarray(1.0 if k_scale is None else k_scale, dtype=np.float32).reshape(-1), (Hkv,))
    vs = np.broadcast_to(np.asarray(1.0 if v_scale is None else v_scale, dtype=np.float32).reshape(-1), (Hkv,))
    out = np.zeros((B, Hq, D), dtype=np.float32)
    lse = np.full((B, Hq), -np.inf, dtype=np.float32)
    for b in range(B):
        L = int(seq_lens[b])
        if L <= 0:
            continue
        for h in range(Hkv):
            K = _dequant(_gather(k_cache, block_table[b], L, h, bs), ks[h], fp8, act)
            V = _dequant(_gather(v_cache, block_table[b], L, h, bs), vs[h], fp8, act)
            acc, M, Ls = _attend_range(q[b, h * G:(h + 1) * G], K, V, sm_scale, act, tile, 0, L)
            out[b, h * G:(h + 1) * G] = acc / Ls[:, None]                   # :606
            lse[b, h * G:(h + 1) * G] = M + np.log(Ls)
    out = act_round(out, act)
    return (out, lse) if return_lse else out


def paged_decode_split_ref(q, k_cache, v_cache, block_table, seq_lens, sm_scale, k_scale=None, v_scale=None, fp8=False,
                           act="bf16", tile=16, num_segments=16):
    """The 3D (split-KV) kernel + reduce_segments: every sequence is cut into `num_segments` runs of
    tiles_per_segment = cdiv(L, num_segments*tile) tiles (:297-300); partials (acc, M, L) are merged by
    reduce_segments (:690-733)."""
    B, Hq, D = q.shape
    bs, Hkv = k_cache.shape[1], k_cache.shape[2]
    G = Hq // Hkv
    ks = np.broadcast_to(np.asarray(1.0 if k_scale is None else k_scale, dtype=np.float32).reshape(-1), (Hkv,))
    vs = np.broadcast_to(np.asarray(1.0 if v_scale is None else v_scale, dtype=np.float32).reshape(-1), (Hkv,))
    out = np.zeros((B, Hq, D), dtype=np.float32)
    for b in range(B):
        L = int(seq_lens[b])
        if L <= 0:
            continue
        tps = -(-L // (num_segments * tile))
        for h in range(Hkv):
            K = _dequant(_gather(k_cache, block_table[b], L, h, bs), ks[h], fp8, act)
            V = _dequant(_gather(v_cache, block_table[b], L, h, bs), vs[h], fp8, act)
            accs, Ms, Ls = [], [], []
            for sgm in range(num_segments):
                lo = sgm * tps * tile
                if lo >= L:
                    break
                a, M, l = _attend_range(q[b, h * G:(h + 1) * G], K, V, sm_scale, act, tile, lo, min(L, lo + tps * tile))
                accs.append(a); Ms.append(M); Ls.append(l)
            out[b, h * G:(h + 1) * G] = merge_partials_ref(np.stack(accs), np.stack(Ms), np.stack(Ls))
    return act_round(out, act)


def merge_partials_ref(acc, M, Lsum):
    """reduce_segments (:690-733): acc [S, G, D], M/L [S, G] -> [G, D]."""
    overall_max = M.max(axis=0)
    w = np.exp(M - overall_max[None]).astype(np.float32)
    expsum = (Lsum * w).sum(axis=0, dtype=np.float32)
    num = (acc * w[:, :, None]).sum(axis=0, dtype=np.float32)
    return np.where(expsum[:, None] == 0.0, np.float32(0.0), num / expsum[:, None]).astype(np.float32)


def merge_lse_ref(parts_o, parts_lse):
    """merge_attn_states generalised to P parts: parts_o [P, n, Hq, D], parts_lse [P, n, Hq] (natural log)."""
    mx = parts_lse.max(axis=0)
    with np.errstate(invalid="ignore"):   # -inf - -inf for rows that are empty on every part
        w = np.where(np.isneginf(parts_lse), 0.0, np.exp(parts_lse - mx[None])).astype(np.float32)
    den = w.sum(axis=0, dtype=np.float32)
    o = (parts_o * w[..., None]).sum(axis=0, dtype=np.float32) / np.where(den == 0, 1, den)[..., None]
    lse = np.where(den > 0, mx + np.log(np.where(den > 0, den, 1)), -np.inf).astype(np.float32)
    return o.astype(np.float32), lse


def paged_decode_exact(q, k_cache, v_cache, block_table, seq_lens, sm_scale, k_scale=None, v_scale=None, fp8=False):
    """float64 mathematical definition (no intermediate rounding): the yardstick both the reference-style oracle and
    the CUDA kernel are measured against in the accuracy tests."""
    B, Hq, D = q.shape
    bs, Hkv = k_cache.shape[1], k_cache.shape[2]
    G = Hq // Hkv
    ks = np.broadcast_to(np.asarray(1.0 if k_scale is None else k_scale, dtype=np.float64).reshape(-1), (Hkv,))
    vs = np.broadcast_to(np.asarray(1.0 if v_scale is None else v_scale, dtype=np.float64).reshape(-1), (Hkv,))
    out = np.zeros((B, Hq, D), dtype=np.float64)
    for b in range(B):
        L = int(seq_lens[b])
        if L <= 0:
            continue
        for h in range(Hkv):
            Kr, Vr = _gather(k_cache, block_table[b], L, h, bs), _gather(v_cache, block_table[b], L, h, bs)
            K = (e4m3_decode(Kr).astype(np.float64) * ks[h]) if fp8 else Kr.astype(np.float64)
            V = (e4m3_decode(Vr).astype(np.float64) * vs[h]) if fp8 else Vr.astype(np.float64)
            S = sm_scale * (q[b, h * G:(h + 1) * G].astype(np.float64) @ K.T)
            S -= S.max(axis=1, keepdims=True)
            P = np.exp(S)
            out[b, h * G:(h + 1) * G] = (P @ V) / P.sum(axis=1, keepdims=True)
    return out


# ------------------------------------------------------------------------------------------------
# paged prefill / mixed-batch attention (q_len >= 1 per request, causal)
# ------------------------------------------------------------------------------------------------


def paged_prefill_ref(q, query_start_loc, k_cache, v_cache, block_table, seq_lens, sm_scale, k_scale=None, v_scale=None,
                      fp8=False, act="bf16", exact=False, rows=None):
    """kernel_unified_attention with q_len > 1 ($VLLM/v1/attention/ops/triton_unified_attention.py): query token t of request
    r (q_len = qsl[r+1] - qsl[r] tokens, the LAST q_len positions of its seq_len) attends to cache positions
    [0, context_len + t + 1) with context_len = seq_len - q_len (:286-303 and the causal mask :511-520); dequant, softmax and
    P.to(V.dtype) @ V as in paged_decode_ref (one tile spanning the row: the tile size only changes fp32 summation order).
    q float32 [T, Hq, D] (activation-representable) -> float32 [T, Hq, D] rounded to `act`.
    exact=True: float64 with no intermediate rounding and an unrounded result (the accuracy yardstick).
    rows: optional iterable of token indices to evaluate (others stay 0) -- full-size cases sample rows."""
    T, Hq, D = q.shape
    bs, Hkv = k_cache.shape[1], k_cache.shape[2]
    G = Hq // Hkv
    ft = np.float64 if exact else np.float32
    ks = np.broadcast_to(np.asarray(1.0 if k_scale is None else k_scale, dtype=ft).reshape(-1), (Hkv,))
    vs = np.broadcast_to(np.asarray(1.0 if v_scale is None else v_scale, dtype=ft).reshape(-1), (Hkv,))
    out = np.zeros((T, Hq, D), dtype=ft)
    want = None if rows is None else set(int(t) for t in rows)
    for r in range(len(seq_lens)):
        q0, q1 = int(query_start_loc[r]), int(query_start_loc[r + 1])
        L, ql = int(seq_lens[r]), q1 - q0
        if ql <= 0 or L <= 0:
            continue
        toks = [t for t in range(ql) if want is None or (q0 + t) in want]
        if not toks:
            continue
        ctx = L - ql
        Lmax = min(L, ctx + max(toks) + 1)
        lim = np.minimum(L, ctx + np.asarray(toks) + 1)                                  # [n]
        mask = np.arange(Lmax)[None, :] < lim[:, None]                                   # [n, Lmax]
        for h in range(Hkv):
            Kr, Vr = _gather(k_cache, block_table[r], Lmax, h, bs), _gather(v_cache, block_table[r], Lmax, h, bs)
            if exact:
                K = (e4m3_decode(Kr).astype(ft) * ks[h]) if fp8 else Kr.astype(ft)
                V = (e4m3_decode(Vr).astype(ft) * vs[h]) if fp8 else Vr.astype(ft)
            else:
                K, V = _dequant(Kr, ks[h], fp8, act), _dequant(Vr, vs[h], fp8, act)
            for g in range(G):
                hq = h * G + g
                S = ft(sm_scale) * (q[[q0 + t for t in toks], hq].astype(ft) @ K.T)       # [n, Lmax]
                S = np.where(mask, S, -np.inf)
                P = np.exp(S - S.max(axis=1, keepdims=True)).astype(ft)
                den = P.sum(axis=1, keepdims=True, dtype=ft)
                Pv = P if exact else act_round(P, act)
                out[[q0 + t for t in toks], hq] = (Pv @ V).astype(ft) / den
    return out if exact else act_round(out, act)


# ------------------------------------------------------------------------------------------------
# seeded synthetic inputs shared by tests, smoke and bench
# ------------------------------------------------------------------------------------------------


def make_block_table(seq_lens, block_size, seed=99, spare=1.1):
    """Random permutation of physical blocks; unused entries 0 (vLLM clears rows to 0, block_table.py:169-171)."""
    rng = np.random.default_rng(seed)
    need = [-(-int(L) // block_size) for L in seq_lens]
    nb = max(2, int(sum(need) * spare) + 1)
    perm = rng.permutation(nb)
    width = max(1, max(need) if need else 1)
    bt = np.zeros((len(seq_lens), width), dtype=np.int32)
    pos = 0
    for i, n in enumerate(need):
        bt[i, :n] = perm[pos:pos + n]
        pos += n
    return bt, nb


def slots_for_positions(block_table, positions, block_size):
    """slot = block_table[req, pos // bs] * bs + pos % bs  (_compute_slot_mapping_kernel, block_table.py:141-164)."""
    out = np.empty(len(positions), dtype=np.int64)
    for i, (r, p) in enumerate(positions):
        out[i] = -1 if r < 0 else int(block_table[r, p // block_size]) * block_size + p % block_size
    return out
