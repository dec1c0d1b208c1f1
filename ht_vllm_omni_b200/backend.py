"""vLLM attention-backend plugin surface for the B200 paged-decode path.

ht-vllm-omni's AR stages (Qwen3-TTS talker, Qwen3-Omni thinker) reach paged attention only through vLLM's
`AttentionBackend` / `AttentionMetadataBuilder` / `AttentionImpl` triple ($VLLM/v1/attention/backend.py:55-342,
516-663, 763-840; caller: vllm_omni/worker/gpu_ar_model_runner.py:228-298).  This module provides that triple over
`torch.ops.b200attn.*` (ops.py -> libb200attn.so), so the hot path is swapped by *registration*, not by editing the
reference:

    register_backend(AttentionBackendEnum.CUSTOM, "ht_vllm_omni_b200.backend.B200PagedAttentionBackend")

and `attention_backend: CUSTOM` in the stage's engine_args (`plugin.register()` does the call).

What the reference does at each hook, and what we do instead:
  * do_kv_cache_update  -> torch.ops._C_cache_ops.reshape_and_cache_flash ($VLLM/.../flashinfer.py:1828-1855)
                         -> torch.ops.b200attn.kv_write (kernel 1)
  * forward (decode)    -> trtllm_batch_decode_with_kv_cache / unified_attention (flashinfer.py:1751-1822,
                           triton_attn.py:526-672)   -> torch.ops.b200attn.paged_decode (kernel 2 + in-kernel merge)
  * MetadataBuilder.build -> FlashInfer `plan` / Triton segment buffers (flashinfer.py:886-1248, triton_attn.py:216-268)
                         -> torch.ops.b200attn.decode_plan (device-side, sync-free) into buffers owned by the builder

Requests with more than one query token (chunked prefill, spec-decode verification, mixed batches) go to the tcgen05
prefill kernel (torch.ops.b200attn.paged_prefill, FP8 KV) while the one-token requests of the same batch stay on the decode
kernel.  The routing is done ON THE DEVICE: the decode plan gets length 0 for the tokens of multi-token requests, the
prefill plan gets no tile for the one-token requests, and both kernels are launched over the whole batch -- so nothing
about the split is a host-side value baked into a captured launch, and no batch reordering is asked of the model runner.
With a 16-bit KV cache (no tensor-core prefill kernel yet) every query token becomes a decode row with its own causal
length and a `row_map` entry pointing at its request's block-table row, as in round 1.

There is no fallback to another backend or to the CPU: without libb200attn.so / a CUDA device the ops raise.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import ClassVar

import torch

from vllm.v1.attention.backend import (
    AttentionBackend,
    AttentionCGSupport,
    AttentionImpl,
    AttentionLayer,
    AttentionMetadataBuilder,
    AttentionType,
    CommonAttentionMetadata,
    MultipleOf,
)

from . import ops
from ._lib import B200AttnError

BACKEND_PATH = "ht_vllm_omni_b200.backend.B200PagedAttentionBackend"
_FP8_NAMES = ("fp8", "fp8_e4m3")
PER_TOKEN_HEAD = "fp8_per_token_head"   # dynamic per-(token, head) scales in float32 side caches (vLLM >= 0.20, KVQuantMode.FP8_PER_TOKEN_HEAD)


def is_fp8_kv(kv_cache_dtype: str) -> bool:
    return kv_cache_dtype in _FP8_NAMES


def expand_query_rows(query_start_loc: torch.Tensor, seq_lens: torch.Tensor, num_tokens: int):
    """Per-token decode rows for a batch with multi-token requests (no host sync; shapes depend only on num_tokens).

    Token t of request r (tokens query_start_loc[r] .. query_start_loc[r+1]-1, the LAST one at position seq_lens[r]-1)
    attends causally to the first `seq_lens[r] - (query_start_loc[r+1] - 1 - t)` cache positions: vLLM's
    context_len + q_pos + 1 (triton_unified_attention.py:286-303).  Returns (row_seq_lens int32 [T], row_map int32 [T]).
    Tokens past query_start_loc[-1] (CUDA-graph padding) get length 0, which the kernel skips.
    """
    t = torch.arange(num_tokens, device=query_start_loc.device, dtype=query_start_loc.dtype)
    num_reqs = seq_lens.shape[0]
    req = torch.bucketize(t, query_start_loc[1:num_reqs + 1], right=True)       # request of token t; == num_reqs for padding
    pad = req >= num_reqs
    req = req.clamp_max(num_reqs - 1)
    last = query_start_loc[1:num_reqs + 1][req] - 1
    lens = seq_lens[req].to(torch.int64) - (last.to(torch.int64) - t.to(torch.int64))
    lens = torch.where(pad, torch.zeros_like(lens), lens.clamp_min(0))
    return lens.to(torch.int32), req.to(torch.int32)


class B200PagedAttentionBackend(AttentionBackend):
    """Static description of the backend; mirrors FlashInferBackend's answers where the layout is shared
    ($VLLM/v1/attention/backends/flashinfer.py:326-436)."""

    accept_output_buffer: bool = True
    supported_dtypes: ClassVar[list[torch.dtype]] = [torch.float16, torch.bfloat16]
    supported_kv_cache_dtypes: ClassVar[list[str]] = ["auto", "float16", "bfloat16", "fp8", "fp8_e4m3", PER_TOKEN_HEAD]
    # forward() writes the step's new tokens itself: on a pure-decode step the store happens INSIDE the decode kernel
    # (b200attn_decode_kv_write: one launch per layer, no kernel boundary between the write and the read); the reference's
    # runner then hands forward() the unpadded slot_mapping (vllm_omni/worker/gpu_ar_model_runner.py:231-244, the same switch
    # as $VLLM/v1/worker/gpu_model_runner.py:4097-4104).  B200ATTN_SEPARATE_KV_UPDATE=1 restores the two-op contract
    # (unified_kv_cache_update -> do_kv_cache_update, then forward), which is also what vLLM's RoPE + KV-cache fusion pass
    # needs to find (do_rope_and_kv_cache_update below).
    forward_includes_kv_cache_update: bool = os.environ.get("B200ATTN_SEPARATE_KV_UPDATE", "0") != "1"

    @staticmethod
    def get_name() -> str:
        return "CUSTOM"   # Attention.__init__ maps the name back through AttentionBackendEnum[...] (attention.py:386)

    @staticmethod
    def get_impl_cls() -> type["B200PagedAttentionImpl"]:
        return B200PagedAttentionImpl

    @staticmethod
    def get_builder_cls() -> type["B200PagedAttentionMetadataBuilder"]:
        return B200PagedAttentionMetadataBuilder

    @staticmethod
    def get_supported_kernel_block_sizes() -> list[int | MultipleOf]:
        return [MultipleOf(16)]   # the TMA box is 16 tokens x 128 bytes; a page holds a whole number of boxes

    @classmethod
    def get_supported_head_sizes(cls) -> list[int]:
        # 128: the TMA / tcgen05 kernels; 64 and 256: the CUDA-core split-KV kernel (csrc/decode_simt.cu) -- the talker config
        # class has no head_dim field, so class defaults resolve to hidden / heads = 64 (configuration_qwen3_tts.py:376-398)
        return [64, 128, 256]

    @staticmethod
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int, head_size: int,
                           cache_dtype_str: str = "auto") -> tuple[int, ...]:
        # logical shape shared with FlashInfer / Triton / FlashAttention-on-sm100 (flashinfer.py:357-368); the KV
        # hand-off code of the reference indexes it as [block, 0|1, slot, head, dim]
        # (vllm_omni/distributed/omni_connectors/kv_transfer_manager.py:224-301)
        if block_size % 16 != 0:
            raise ValueError("B200 paged attention needs block_size % 16 == 0")
        if cache_dtype_str == PER_TOKEN_HEAD:
            # same byte budget as TritonAttentionBackend (triton_attn.py:312-334: one float32 scale per (slot, head) inline);
            # the bytes of a block are laid out by ops.per_token_head_views (data pages first, scale runs behind them)
            return (num_blocks, 2, block_size, num_kv_heads, head_size + 4)
        return (num_blocks, 2, block_size, num_kv_heads, head_size)

    @staticmethod
    def get_kv_cache_stride_order(include_num_layers_dimension: bool = False) -> tuple[int, ...]:
        # HND physical order: [block][K|V][head][slot][dim] -> one (block, head) K page is bs*D contiguous bytes
        if include_num_layers_dimension:
            return (1, 2, 4, 0, 3, 5)   # (num_blocks, 2, num_kv_heads, num_layers, block_size, head_size)
        return (0, 1, 3, 2, 4)

    @classmethod
    def get_required_kv_cache_layout(cls):
        return "HND"

    @classmethod
    def supports_per_head_quant_scales(cls) -> bool:
        return True

    @classmethod
    def supports_compute_capability(cls, capability) -> bool:
        return getattr(capability, "major", None) == 10   # sm_100a cubins only

    @classmethod
    def supports_attn_type(cls, attn_type: str) -> bool:
        return attn_type == AttentionType.DECODER

    @classmethod
    def supports_combination(cls, head_size, dtype, kv_cache_dtype, block_size, use_mla, has_sink, use_sparse,
                             device_capability) -> str | None:
        if kv_cache_dtype in ("float16", "bfloat16") and dtype is not None:
            want = torch.float16 if kv_cache_dtype == "float16" else torch.bfloat16
            if dtype != want:
                return f"16-bit KV cache ({kv_cache_dtype}) must match the model dtype ({dtype})"
        return None


@dataclass
class B200PagedAttentionMetadata:
    """Per-step, per-KV-cache-group metadata.  All tensors are views of vLLM's persistent buffers or of the builder's
    own pre-allocated buffers, so the object is CUDA-graph safe (capture and replay see the same addresses)."""

    num_actual_tokens: int
    num_rows: int                     # decode rows handed to the kernel (== tokens)
    max_query_len: int
    seq_lens: torch.Tensor            # int32 [num_rows]: causal length of each row
    block_table: torch.Tensor         # int32 [num_reqs, max_blocks]
    row_map: torch.Tensor | None      # int32 [num_rows] block-table row per decode row; None = identity (pure decode)
    slot_mapping: torch.Tensor
    workspace: "ops.DecodeWorkspace"  # plan + split-KV partial buffers (owned by the builder)
    use_cascade: bool = False
    pure_decode: bool = False         # one query token per request and rows == requests (host-known: max_query_len, counts)
    # mixed batch routed between the two kernels (FP8 KV): rows of multi-token requests have seq_lens == 0 above and are
    # served by the prefill kernel from the request-level tensors below
    has_decode: bool = True
    has_prefill: bool = False
    num_reqs: int = 0
    req_seq_lens: torch.Tensor | None = None      # int32 [num_reqs]
    query_start_loc: torch.Tensor | None = None   # int32 [num_reqs + 1]
    tile_prefix: torch.Tensor | None = None       # int32 [num_reqs + 1] (builder-owned)


class B200PagedAttentionMetadataBuilder(AttentionMetadataBuilder[B200PagedAttentionMetadata]):
    """Builds the split-KV plan once per step on the device (kernel `plan_kernel`, no host sync, no allocation)."""

    _cudagraph_support: ClassVar[AttentionCGSupport] = AttentionCGSupport.ALWAYS
    reorder_batch_threshold: int | None = None   # decode / prefill routing happens on the device: no reordering needed
    prefill_min_query_len: ClassVar[int] = 2      # requests with at least this many query tokens go to the prefill kernel

    @classmethod
    def get_cudagraph_support(cls, vllm_config, kv_cache_spec) -> AttentionCGSupport:
        """Two-op contract: every launch reads the builder's fixed buffers, so one FULL graph serves any batch (ALWAYS).
        When forward() owns the KV write, a pure-decode step takes the in-kernel-write launch, which is only valid for one
        token per request -- a host-side choice that a captured graph would freeze.  FULL graphs are therefore offered for
        UNIFORM batches only (every request the same query length: plain decode, or 1 + k spec-decode tokens); mixed batches
        run the hooks eagerly between piecewise graphs."""
        if B200PagedAttentionBackend.forward_includes_kv_cache_update:
            # uniform batches: q_len == 1 decodes take the in-kernel-write launch, q_len == 1 + k (spec-decode / the talker's
            # MTP verification, vllm_omni/worker/gpu_model_runner.py:1266-1303) the general one; a graph captured for one
            # uniform query length is only ever replayed for that length, so the host-side choice is graph-key invariant
            return AttentionCGSupport.UNIFORM_BATCH
        return cls._cudagraph_support

    def __init__(self, kv_cache_spec, layer_names, vllm_config, device):
        super().__init__(kv_cache_spec, layer_names, vllm_config, device)
        model_config = vllm_config.model_config
        parallel_config = vllm_config.parallel_config
        self.block_size = kv_cache_spec.block_size
        self.num_heads_q = model_config.get_num_attention_heads(parallel_config)
        self.num_heads_kv = model_config.get_num_kv_heads(parallel_config)
        self.head_dim = model_config.get_head_size()
        kv_dtype = getattr(kv_cache_spec, "dtype", torch.bfloat16)
        self.kv_torch_dtype = torch.uint8 if kv_dtype in (torch.uint8, torch.float8_e4m3fn) else kv_dtype
        # rows = tokens: size every buffer for the largest batch vLLM can schedule
        self.max_rows = int(vllm_config.scheduler_config.max_num_batched_tokens)
        self.workspace = ops.DecodeWorkspace(device, self.max_rows, self.kv_torch_dtype, head_dim=self.head_dim,
                                             group_size=max(1, self.num_heads_q // max(1, self.num_heads_kv)))
        self._row_lens = torch.zeros(self.max_rows, dtype=torch.int32, device=device)
        self._row_map = torch.zeros(self.max_rows, dtype=torch.int32, device=device)
        self._identity = torch.arange(self.max_rows, dtype=torch.int32, device=device)
        self.max_reqs = int(getattr(vllm_config.scheduler_config, "max_num_seqs", None) or self.max_rows)
        self._tile_prefix = torch.zeros(self.max_reqs + 1, dtype=torch.int32, device=device)
        # the tensor-core prefill kernel reads the FP8 cache with static scales; 16-bit caches and per-token-head scales keep the
        # per-token decode rows
        mode = getattr(kv_cache_spec, "kv_quant_mode", None)
        self.per_token_head = bool(getattr(mode, "is_per_token_head", False)) or \
            getattr(getattr(vllm_config, "cache_config", None), "cache_dtype", None) == PER_TOKEN_HEAD
        self.use_prefill_kernel = self.kv_torch_dtype == torch.uint8 and self.head_dim == 128 and not self.per_token_head
        self._capturing = False

    def build(self, common_prefix_len: int, common_attn_metadata: CommonAttentionMetadata,
              fast_build: bool = False) -> B200PagedAttentionMetadata:
        """Runs eagerly before every step (and before every CUDA-graph replay).  The kernels a FULL graph captured read
        their row lengths and row map from the builder's OWN fixed buffers, whatever the batch looks like: a pure-decode
        batch fills them with (seq_lens, identity), a batch with multi-token requests with the expanded rows.  Which of
        the two it is therefore never changes a pointer baked into a captured launch."""
        m = common_attn_metadata
        T = m.num_actual_tokens
        if T > self.max_rows:
            raise B200AttnError(f"{T} tokens exceed the planned maximum of {self.max_rows}")
        R = m.num_reqs
        has_decode, has_prefill = True, False
        pure = m.max_query_len <= 1 and T == R
        if pure:
            self._row_lens[:T].copy_(m.seq_lens[:T])                # pure decode: rows == requests
            self._row_map[:T].copy_(self._identity[:T])
        else:
            lens, rmap = expand_query_rows(m.query_start_loc, m.seq_lens, T)
            if self.use_prefill_kernel and R <= self.max_reqs:
                qsl = m.query_start_loc[:R + 1]
                short = (qsl[1:] - qsl[:-1]) < self.prefill_min_query_len                 # [R] on the device
                lens = torch.where(short[rmap.long()], lens, torch.zeros_like(lens))       # the prefill kernel's tokens: no decode row
                torch.ops.b200attn.prefill_plan(qsl, self._tile_prefix, self.num_heads_q, self.num_heads_kv, self.prefill_min_query_len)
                has_prefill = True
                if not self._capturing and m.query_start_loc_cpu is not None:
                    # eager step: skip the launch of a kernel that has no request (host copy of the offsets, no sync)
                    ql = m.query_start_loc_cpu[1:R + 1] - m.query_start_loc_cpu[:R]
                    has_decode = bool((ql < self.prefill_min_query_len).any())
                    has_prefill = bool((ql >= self.prefill_min_query_len).any())
            self._row_lens[:T].copy_(lens)
            self._row_map[:T].copy_(rmap)
        seq_lens, row_map = self._row_lens[:T], self._row_map[:T]
        if has_decode:
            self.workspace.make_plan(seq_lens, self.num_heads_q, self.num_heads_kv, m.block_table_tensor, self.block_size, row_map)
        return B200PagedAttentionMetadata(
            num_actual_tokens=T, num_rows=T, max_query_len=m.max_query_len, seq_lens=seq_lens,
            block_table=m.block_table_tensor, row_map=row_map, slot_mapping=m.slot_mapping, workspace=self.workspace,
            pure_decode=pure, has_decode=has_decode, has_prefill=has_prefill, num_reqs=R, req_seq_lens=m.seq_lens[:R],
            query_start_loc=m.query_start_loc[:R + 1], tile_prefix=self._tile_prefix[:R + 1])

    def build_for_cudagraph_capture(self, common_attn_metadata: CommonAttentionMetadata) -> B200PagedAttentionMetadata:
        # the plan kernels read seq_lens / query_start_loc on the device at replay time; nothing about the batch is baked
        # in at capture, and a captured mixed batch launches BOTH kernels (either may find no work at replay)
        self._capturing = True
        try:
            return self.build(0, common_attn_metadata)
        finally:
            self._capturing = False

    def use_cascade_attention(self, *args, **kwargs) -> bool:
        return False


class B200PagedAttentionImpl(AttentionImpl[B200PagedAttentionMetadata]):
    # vLLM's decode-context-parallel gate ($VLLM/v1/worker/cp_utils.py:31) asserts on this flag and would then call forward()
    # with rank-local block tables and GLOBAL seq_lens.  forward() does not run that flow (query all-gather, local lengths,
    # LSE combine), so the flag stays off and __init__ refuses a DCP group; the sequence-split exchange of this repo is
    # driven explicitly through forward_with_lse + parallel.PeerExchange / dcp_combine instead.
    can_return_lse_for_decode: bool = False

    def __init__(self, num_heads: int, head_size: int, scale: float, num_kv_heads: int | None = None,
                 alibi_slopes: list[float] | None = None, sliding_window: int | None = None,
                 kv_cache_dtype: str = "auto", logits_soft_cap: float | None = None,
                 attn_type: str = AttentionType.DECODER, kv_sharing_target_layer_name: str | None = None,
                 sinks: torch.Tensor | None = None, **extra) -> None:
        # unsupported features are rejected at construction (selection time), never on the hot path
        if alibi_slopes is not None:
            raise NotImplementedError("B200 paged attention: ALiBi is not supported")
        if sliding_window is not None:
            raise NotImplementedError("B200 paged attention: sliding window is not supported")
        if logits_soft_cap:
            raise NotImplementedError("B200 paged attention: logits soft cap is not supported")
        if sinks is not None:
            raise NotImplementedError("B200 paged attention: attention sinks are not supported")
        if attn_type != AttentionType.DECODER:
            raise NotImplementedError("B200 paged attention: decoder self-attention only")
        if head_size not in (64, 128, 256):
            raise NotImplementedError(f"B200 paged attention: head_size {head_size} unsupported (64, 128 or 256)")
        if kv_cache_dtype == PER_TOKEN_HEAD and head_size != 128:
            raise NotImplementedError("B200 paged attention: fp8_per_token_head needs head_size 128")
        if getattr(self, "dcp_world_size", 1) > 1:   # set by AttentionImplBase.__new__ ($VLLM/v1/attention/backend.py:731-757)
            raise NotImplementedError("B200 paged attention: vLLM decode-context-parallel groups are not supported")
        if kv_cache_dtype not in B200PagedAttentionBackend.supported_kv_cache_dtypes:
            raise NotImplementedError(f"B200 paged attention: kv_cache_dtype {kv_cache_dtype!r} unsupported")
        self.num_heads = num_heads
        self.head_size = head_size
        self.scale = float(scale)
        self.num_kv_heads = num_kv_heads if num_kv_heads is not None else num_heads
        if self.num_heads % self.num_kv_heads != 0:
            raise ValueError("num_heads must be a multiple of num_kv_heads")
        self.num_queries_per_kv = self.num_heads // self.num_kv_heads
        self.kv_cache_dtype = kv_cache_dtype
        self.attn_type = attn_type
        self.kv_sharing_target_layer_name = kv_sharing_target_layer_name
        self.alibi_slopes = None
        self.sliding_window = (-1, -1)
        self.logits_soft_cap = 0.0
        self.supports_quant_query_input = False   # Q stays bf16/fp16: K is expanded exactly in-kernel instead
        self._unit_scale: torch.Tensor | None = None
        self.per_token_head = kv_cache_dtype == PER_TOKEN_HEAD
        self._pth_views: dict = {}

    def _per_token_head(self, kv_cache: torch.Tensor):
        """(k, v, k_scale_cache, v_scale_cache) views of a per-token-head cache, built once per cache tensor."""
        key = (kv_cache.data_ptr(), tuple(kv_cache.shape))
        got = self._pth_views.get(key)
        if got is None:
            got = self._pth_views[key] = ops.per_token_head_views(kv_cache.view(torch.uint8) if kv_cache.dtype != torch.uint8 else kv_cache,
                                                                  self.num_kv_heads, self.head_size)
        return got

    # ---- helpers ----
    def _planes(self, kv_cache: torch.Tensor):
        """[num_blocks, 2, block_size, Hkv, D] (any strides) -> K and V planes [num_blocks, block_size, Hkv, D]."""
        if kv_cache.dim() != 5 or kv_cache.shape[1] != 2:
            raise B200AttnError(f"kv_cache must be [num_blocks, 2, block_size, Hkv, D], got {tuple(kv_cache.shape)}")
        k, v = kv_cache.unbind(1)
        if is_fp8_kv(self.kv_cache_dtype) and k.dtype not in (torch.uint8, torch.float8_e4m3fn):
            k, v = k.view(torch.uint8), v.view(torch.uint8)
        return k, v

    def _scales(self, layer: AttentionLayer, device):
        if is_fp8_kv(self.kv_cache_dtype):
            return layer._k_scale, layer._v_scale     # fp32 device tensors, 1 or Hkv elements (attention.py:94-174)
        if self._unit_scale is None or self._unit_scale.device != device:
            self._unit_scale = torch.ones(1, dtype=torch.float32, device=device)
        return self._unit_scale, self._unit_scale

    # ---- hooks ----
    def do_kv_cache_update(self, layer: AttentionLayer, key: torch.Tensor, value: torch.Tensor,
                           kv_cache: torch.Tensor, slot_mapping: torch.Tensor) -> None:
        """Replaces reshape_and_cache_flash.  key/value are [T, Hkv, D] slices of the fused qkv activation (strided);
        slot_mapping may be longer than the real token count under padding: -1 entries are skipped in-kernel."""
        if self.kv_sharing_target_layer_name is not None:
            return   # the target layer already wrote this cache (attention.py:700-706 semantics)
        if self.per_token_head:
            k_cache, v_cache, ksc, vsc = self._per_token_head(kv_cache)
            n = min(slot_mapping.shape[0], key.shape[0])
            torch.ops.b200attn.kv_write_per_token_head(key[:n].view(n, self.num_kv_heads, self.head_size),
                                                       value[:n].view(n, self.num_kv_heads, self.head_size), k_cache, v_cache, ksc, vsc,
                                                       slot_mapping[:n])
            return
        k_cache, v_cache = self._planes(kv_cache)
        k_scale, v_scale = self._scales(layer, key.device)
        n = min(slot_mapping.shape[0], key.shape[0])
        torch.ops.b200attn.kv_write(key[:n].view(n, self.num_kv_heads, self.head_size),
                                    value[:n].view(n, self.num_kv_heads, self.head_size),
                                    k_cache, v_cache, slot_mapping[:n], k_scale, v_scale)

    def fused_rope_kvcache_supported(self) -> bool:
        """vLLM's RoPE+KV-cache fusion pass ($VLLM/compilation/passes/fusion/rope_kvcache_fusion.py:50) asks the impl; ours
        does both in one launch (csrc/rope_kv_write.cu).  KV-sharing layers keep the unfused path (they write no cache).
        The pass rewrites rotary_embedding + unified_kv_cache_update, an op that only exists under the two-op contract."""
        return (self.kv_sharing_target_layer_name is None and not self.per_token_head
                and not B200PagedAttentionBackend.forward_includes_kv_cache_update)

    def do_rope_and_kv_cache_update(self, layer: AttentionLayer, query: torch.Tensor, key: torch.Tensor, value: torch.Tensor,
                                    positions: torch.Tensor, cos_sin_cache: torch.Tensor, is_neox: bool,
                                    kv_cache: torch.Tensor, layer_slot_mapping: torch.Tensor) -> None:
        """Replaces RotaryEmbedding.forward + reshape_and_cache_flash ($VLLM/v1/attention/backend.py:823-840; the ROCm
        reference body is triton_attn.py:770-804).  query/key are rotated in place; the rotated key and the value go to the
        cache.  cos_sin_cache must already be in the activation dtype (RotaryEmbedding._match_cos_sin_cache_dtype)."""
        k_cache, v_cache = self._planes(kv_cache)
        k_scale, v_scale = self._scales(layer, key.device)
        n = min(layer_slot_mapping.shape[0], key.shape[0])
        if cos_sin_cache.dtype != query.dtype:
            cos_sin_cache = cos_sin_cache.to(query.dtype)
        torch.ops.b200attn.rope_kv_write(query[:n].view(n, self.num_heads, self.head_size),
                                         key[:n].view(n, self.num_kv_heads, self.head_size),
                                         value[:n].view(n, self.num_kv_heads, self.head_size),
                                         positions[:n].contiguous(), cos_sin_cache, bool(is_neox), k_cache, v_cache,
                                         layer_slot_mapping[:n], k_scale, v_scale)

    def do_norm_rope_and_kv_cache_update(self, layer: AttentionLayer, query: torch.Tensor, key: torch.Tensor, value: torch.Tensor,
                                         positions: torch.Tensor, rotary_emb, q_norm, k_norm, kv_cache: torch.Tensor,
                                         layer_slot_mapping: torch.Tensor) -> None:
        """The Qwen3 attention prologue in one launch: q_norm / k_norm (per-head RMSNorm), the rotary embedding -- plain, or
        M-RoPE when `positions` is [3, T] (the Qwen3-Omni thinker, qwen3_omni_moe_thinker.py:1633-1728) -- and the KV write
        ($VLLM/model_executor/models/qwen3.py:153-160 followed by unified_kv_cache_update).  `rotary_emb` is the layer's
        RotaryEmbedding / MRotaryEmbedding module (cos_sin_cache, is_neox_style, mrope_section, mrope_interleaved are read
        from it), `q_norm` / `k_norm` its RMSNorm modules or None.  query / key are updated in place."""
        k_cache, v_cache = self._planes(kv_cache)
        k_scale, v_scale = self._scales(layer, key.device)
        n = min(layer_slot_mapping.shape[0], key.shape[0])
        cache = rotary_emb.cos_sin_cache
        if cache.dtype != query.dtype or cache.device != query.device:
            cache = cache.to(device=query.device, dtype=query.dtype)
        empty = query.new_empty(0)
        qw = q_norm.weight.data.to(query.dtype) if q_norm is not None else empty
        kw = k_norm.weight.data.to(query.dtype) if k_norm is not None else empty
        eps = float(getattr(q_norm if q_norm is not None else k_norm, "variance_epsilon", 0.0)) if (q_norm is not None or k_norm is not None) else 0.0
        pos = positions[..., :n].contiguous()
        section = list(getattr(rotary_emb, "mrope_section", None) or [0, 0, 0])
        torch.ops.b200attn.norm_rope_kv_write(query[:n].view(n, self.num_heads, self.head_size),
                                              key[:n].view(n, self.num_kv_heads, self.head_size),
                                              value[:n].view(n, self.num_kv_heads, self.head_size), pos, cache,
                                              bool(rotary_emb.is_neox_style), qw, kw, eps, section,
                                              bool(getattr(rotary_emb, "mrope_interleaved", False)), k_cache, v_cache,
                                              layer_slot_mapping[:n], k_scale, v_scale)

    def forward(self, layer: AttentionLayer, query: torch.Tensor, key: torch.Tensor, value: torch.Tensor,
                kv_cache: torch.Tensor, attn_metadata: B200PagedAttentionMetadata, output: torch.Tensor,
                output_scale: torch.Tensor | None = None, output_block_scale: torch.Tensor | None = None) -> torch.Tensor:
        assert output is not None, "output tensor must be provided (accept_output_buffer)"
        if output_scale is not None or output_block_scale is not None:
            raise NotImplementedError("B200 paged attention: fused output quantisation is not supported")
        if attn_metadata is None:
            return output.fill_(0)    # profiling run (triton_attn.py:557-559)
        md = attn_metadata
        T = md.num_actual_tokens
        if self.per_token_head:
            # dynamic per-(token, head) scales: HMMA consumers with the scales applied to the logits / probabilities; every
            # query token is a decode row (row_map), as for 16-bit caches
            k_cache, v_cache, ksc, vsc = self._per_token_head(kv_cache)
            q = query[:T].view(T, self.num_heads, self.head_size)
            out = output[:T].view(T, self.num_heads, self.head_size)
            if (B200PagedAttentionBackend.forward_includes_kv_cache_update and key is not None and value is not None
                    and self.kv_sharing_target_layer_name is None):
                torch.ops.b200attn.kv_write_per_token_head(key[:T].view(T, self.num_kv_heads, self.head_size),
                                                           value[:T].view(T, self.num_kv_heads, self.head_size), k_cache, v_cache,
                                                           ksc, vsc, md.slot_mapping[:T])
            md.workspace.decode_per_token_head(q, k_cache, v_cache, ksc, vsc, md.block_table, md.seq_lens, out, self.scale,
                                               row_map=md.row_map, fresh_tokens=max(1, md.max_query_len))
            return output
        k_cache, v_cache = self._planes(kv_cache)
        k_scale, v_scale = self._scales(layer, query.device)
        q = query[:T].view(T, self.num_heads, self.head_size)
        out = output[:T].view(T, self.num_heads, self.head_size)
        if (B200PagedAttentionBackend.forward_includes_kv_cache_update and key is not None and value is not None
                and self.kv_sharing_target_layer_name is None):
            # this backend's forward() owns the KV write (attention.py:493 skips unified_kv_cache_update for it)
            k_new = key[:T].view(T, self.num_kv_heads, self.head_size)
            v_new = value[:T].view(T, self.num_kv_heads, self.head_size)
            if md.pure_decode:
                # one token per request: the decode launch stores each row's token right before it reads the row's last tile
                md.workspace.decode_kv_write(q, k_new, v_new, md.slot_mapping[:T], k_cache, v_cache, md.block_table, md.seq_lens,
                                             out, k_scale, v_scale, self.scale, row_map=None, fresh_tokens=1, fused_merge=True)
                return output
            torch.ops.b200attn.kv_write(k_new, v_new, k_cache, v_cache, md.slot_mapping[:T], k_scale, v_scale)
        if md.has_decode:
            # fresh_tokens: the KV write that precedes us on the stream touched at most the last max_query_len positions of
            # a decode row's sequence (1 when the multi-token requests are the prefill kernel's)
            fresh = 1 if md.has_prefill else max(1, md.max_query_len)
            md.workspace.decode(q, k_cache, v_cache, md.block_table, md.seq_lens, out, k_scale, v_scale, self.scale,
                                row_map=md.row_map, fresh_tokens=fresh, fused_merge=True)
        if md.has_prefill:
            torch.ops.b200attn.paged_prefill(q, k_cache, v_cache, md.block_table, md.req_seq_lens, md.query_start_loc,
                                             md.tile_prefix, out, k_scale, v_scale, self.scale, 0)
        return output

    def forward_with_lse(self, layer, query, kv_cache, attn_metadata, output, lse) -> None:
        """Decode returning the natural-log LSE per (row, head): the per-rank piece of a sequence-split (DCP) exchange,
        combined across ranks by torch.ops.b200attn.merge_lse (replaces cp_lse_ag_out_rs, flashinfer.py:1714-1738)."""
        md = attn_metadata
        if md.has_prefill:
            raise NotImplementedError("forward_with_lse serves decode rows only (the prefill kernel returns no LSE)")
        T = md.num_actual_tokens
        k_cache, v_cache = self._planes(kv_cache)
        k_scale, v_scale = self._scales(layer, query.device)
        # rows that hold no token on this rank come back as (o = 0, lse = -inf) from the kernel itself (decode_tc.cu /
        # decode.cu write them when an LSE buffer is given), so a stale buffer can never leak into the combine
        md.workspace.decode(query[:T].view(T, self.num_heads, self.head_size), k_cache, v_cache, md.block_table,
                            md.seq_lens, output[:T].view(T, self.num_heads, self.head_size), k_scale, v_scale,
                            self.scale, lse=lse, row_map=md.row_map, fresh_tokens=max(1, md.max_query_len))
