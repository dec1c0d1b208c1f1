"""Comparison arms: the reference's OWN Blackwell kernels for this path, timed beside ours on the same inputs with the
same method (CUDA graph of one launch per layer cache, CUDA events on the launching stream, L2-cold because the layer
caches together exceed the L2).  Never imported by the package; bench.py and the profiling tools call it.

  flashinfer  flashinfer.decode.trtllm_batch_decode_with_kv_cache -- the TRT-LLM-gen sm100 cubins vLLM's FLASHINFER backend
              runs on capability 10 ($VLLM/v1/attention/backends/flashinfer.py:1751-1822).  Two query dtypes: bf16 (like for
              like with ours) and fp8 (what vLLM does by default for an fp8 cache: supports_quant_query_input,
              flashinfer.py:1317-1323; the query quantisation itself is not timed).
  triton      vllm.v1.attention.ops.triton_unified_attention.unified_attention (TRITON_ATTN backend,
              $VLLM/v1/attention/backends/triton_attn.py:638-670), 2D or 3D split as vLLM's own heuristic picks.
  fa4         vllm.vllm_flash_attn.flash_attn_varlen_func(fa_version=4) for bf16 KV ($VLLM/v1/attention/backends/fa_utils.py:81-83).
  kv_write    torch.ops._C_cache_ops.reshape_and_cache_flash (the compiled CUDA op) for the write row.

Every arm returns {"us_per_layer": ..., "gbs": ...} or {"error": "..."}; an arm that cannot run on the box never fails the bench.
"""
from __future__ import annotations

import math
import os

import torch


def _graph_time(body, layers: int, iters: int = 10, warm: int = 3) -> float:
    """us per layer of `body()` (which launches `layers` kernels) replayed as one CUDA graph."""
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        body()
        body()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    try:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            body()
        run = g.replay
    except Exception:  # noqa: BLE001  (an arm that cannot be captured is timed eagerly; said in the result)
        torch.cuda.synchronize()
        run = body
    for _ in range(warm):
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        run()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / iters / layers, run is not body


def _res(us, graphed, nbytes, extra=None):
    r = {"us_per_layer": us, "gbs": nbytes / us / 1e3, "cuda_graph": bool(graphed)}
    if extra:
        r.update(extra)
    return r


def time_flashinfer(phys_caches, q, bt, lens, max_len, sm_scale, k_scale, v_scale, nbytes, q_fp8: bool):
    """phys_caches: list of [nb, 2, Hkv, bs, D] (HND physical) uint8 / bf16 tensors; q [NL, B, Hq, D] bf16."""
    try:
        from flashinfer.decode import trtllm_batch_decode_with_kv_cache
    except Exception as e:  # noqa: BLE001
        return {"error": f"import flashinfer: {type(e).__name__}: {e}"}
    try:
        fp8 = phys_caches[0].dtype in (torch.uint8, torch.float8_e4m3fn)
        if q_fp8 and not fp8:
            return {"error": "fp8 query only with an fp8 cache"}
        caches = [c.view(torch.float8_e4m3fn) if fp8 else c for c in phys_caches]
        dev = q.device
        ws = torch.zeros(int(os.environ.get("VLLM_FLASHINFER_WORKSPACE_BUFFER_SIZE", 394 * 1024 * 1024)), dtype=torch.uint8, device=dev)
        NL, B, Hq, D = q.shape
        out = torch.zeros(NL, B, Hq, D, dtype=torch.bfloat16, device=dev)
        ks, vs = float(k_scale), float(v_scale)
        q_scale = 1.0
        qq = q
        if q_fp8:
            q_scale = float(q.float().abs().max()) / 448.0
            qq = (q.float() / q_scale).clamp(-448, 448).to(torch.float8_e4m3fn)
        bmm1 = sm_scale * q_scale * (ks if fp8 else 1.0)
        bmm2 = vs if fp8 else 1.0
        bt_c, lens_c = bt.contiguous(), lens.contiguous()

        def body():
            for l in range(NL):
                trtllm_batch_decode_with_kv_cache(query=qq[l], kv_cache=caches[l], workspace_buffer=ws, block_tables=bt_c,
                                                  seq_lens=lens_c, max_seq_len=int(max_len), bmm1_scale=bmm1, bmm2_scale=bmm2,
                                                  out=out[l], kv_layout="HND")

        us, graphed = _graph_time(body, NL)
        return _res(us, graphed, nbytes, {"q_dtype": "fp8_e4m3" if q_fp8 else "bf16", "out": out})
    except Exception as e:  # noqa: BLE001
        torch.cuda.synchronize()
        return {"error": f"{type(e).__name__}: {str(e)[:300]}"}


def time_triton(caches, q, bt, lens, max_len, sm_scale, k_scale_t, v_scale_t, nbytes, Hkv):
    """caches: list of (k_plane, v_plane) [nb, bs, Hkv, D] views (any strides)."""
    try:
        from vllm.v1.attention.ops.triton_unified_attention import unified_attention
        from vllm.v1.kv_cache_interface import KVQuantMode
    except Exception as e:  # noqa: BLE001
        return {"error": f"import vllm triton: {type(e).__name__}: {e}"}
    try:
        fp8 = caches[0][0].dtype in (torch.uint8, torch.float8_e4m3fn)
        NL, B, Hq, D = q.shape
        dev = q.device
        out = torch.zeros(NL, B, Hq, D, dtype=torch.bfloat16, device=dev)
        cu = torch.arange(B + 1, dtype=torch.int32, device=dev)
        kw = {}
        # vLLM's own heuristic ($VLLM/v1/attention/backends/triton_attn.py:54-55,168-183): 3D split-KV with 16 segments when
        # num_seqs <= 128 // Hkv
        segs = 16
        thr = 128 // Hkv
        if B <= thr:
            kw = dict(seq_threshold_3D=thr, num_par_softmax_segments=segs,
                      softmax_segm_output=torch.zeros(B, Hq, segs, D, dtype=torch.float32, device=dev),
                      softmax_segm_max=torch.zeros(B, Hq, segs, dtype=torch.float32, device=dev),
                      softmax_segm_expsum=torch.zeros(B, Hq, segs, dtype=torch.float32, device=dev))
        kd = k_scale_t.reshape(1, -1).expand(B, Hkv) if fp8 else None
        vd = v_scale_t.reshape(1, -1).expand(B, Hkv) if fp8 else None
        cc = [(k.view(torch.float8_e4m3fn), v.view(torch.float8_e4m3fn)) if fp8 else (k, v) for k, v in caches]

        def body():
            for l in range(NL):
                unified_attention(q[l], cc[l][0], cc[l][1], out[l], cu, 1, lens, int(max_len), sm_scale, True, (-1, -1), bt, 0.0,
                                  None, kd, vd, kv_quant_mode=KVQuantMode.FP8_PER_TENSOR if fp8 else KVQuantMode.NONE, **kw)

        us, graphed = _graph_time(body, NL)
        return _res(us, graphed, nbytes, {"split": "3D/16 segments" if kw else "2D", "out": out})
    except Exception as e:  # noqa: BLE001
        torch.cuda.synchronize()
        return {"error": f"{type(e).__name__}: {str(e)[:300]}"}


def time_fa4(caches, q, bt, lens, max_len, sm_scale, nbytes):
    """bf16 KV only: FlashAttention-4 (CuTe-DSL sm100) through vLLM's wrapper, paged KV by block_table."""
    try:
        from vllm.vllm_flash_attn import flash_attn_varlen_func
    except Exception as e:  # noqa: BLE001
        return {"error": f"import vllm_flash_attn: {type(e).__name__}: {e}"}
    try:
        NL, B, Hq, D = q.shape
        dev = q.device
        out = torch.zeros(NL, B, Hq, D, dtype=torch.bfloat16, device=dev)
        cu = torch.arange(B + 1, dtype=torch.int32, device=dev)

        def body():
            for l in range(NL):
                k, v = caches[l]
                flash_attn_varlen_func(q=q[l], k=k, v=v, out=out[l], cu_seqlens_q=cu, max_seqlen_q=1, seqused_k=lens,
                                       max_seqlen_k=int(max_len), softmax_scale=sm_scale, causal=True, block_table=bt,
                                       fa_version=4)

        us, graphed = _graph_time(body, NL)
        return _res(us, graphed, nbytes, {"out": out})
    except Exception as e:  # noqa: BLE001
        torch.cuda.synchronize()
        return {"error": f"{type(e).__name__}: {str(e)[:300]}"}


def time_vllm_kv_write(caches, kn, vn, slots, k_scale_t, v_scale_t, nbytes):
    try:
        import vllm._custom_ops as cops
    except Exception as e:  # noqa: BLE001
        return {"error": f"import vllm._custom_ops: {type(e).__name__}: {e}"}
    try:
        fp8 = caches[0][0].dtype in (torch.uint8, torch.float8_e4m3fn)
        NL = kn.shape[0]
        cc = [(k.view(torch.float8_e4m3fn), v.view(torch.float8_e4m3fn)) if fp8 else (k, v) for k, v in caches]
        ks1, vs1 = k_scale_t[:1].contiguous(), v_scale_t[:1].contiguous()

        def body():
            for l in range(NL):
                cops.reshape_and_cache_flash(kn[l], vn[l], cc[l][0], cc[l][1], slots, "fp8" if fp8 else "auto", ks1, vs1)

        us, graphed = _graph_time(body, NL)
        return _res(us, graphed, nbytes)
    except Exception as e:  # noqa: BLE001
        torch.cuda.synchronize()
        return {"error": f"{type(e).__name__}: {str(e)[:300]}"}


def time_vllm_query_quant(q):
    """The static per-tensor FP8 quantisation of the query that vLLM's Attention.forward issues in front of an FP8-query
    TRT-LLM-gen call (`query_quant`, $VLLM/model_executor/layers/attention/attention.py:461-471; under torch.compile it can
    fuse into the producer of q, eager / custom-op mode launches it as timed here).  q [NL, B, Hq, D] bf16."""
    try:
        import vllm._custom_ops as cops
    except Exception as e:  # noqa: BLE001
        return {"error": f"import vllm._custom_ops: {type(e).__name__}: {e}"}
    try:
        NL, B, Hq, D = q.shape
        scale = torch.full((1,), float(q.float().abs().max()) / 448.0, dtype=torch.float32, device=q.device)
        q2 = q.reshape(NL, B, Hq * D)

        def body():
            for l in range(NL):
                cops.scaled_fp8_quant(q2[l], scale)

        us, graphed = _graph_time(body, NL)
        return _res(us, graphed, B * Hq * D * 3)
    except Exception as e:  # noqa: BLE001
        torch.cuda.synchronize()
        return {"error": f"{type(e).__name__}: {str(e)[:300]}"}


def max_abs_vs(ours: torch.Tensor, other) -> float | None:
    if not isinstance(other, dict) or "out" not in other:
        return None
    o = other.pop("out")
    return float((ours.float() - o.float()).abs().max())


def compare_all(phys_caches, q, bt, lens, sm_scale, k_scale_t, v_scale_t, dec_bytes, wr_bytes, kn=None, vn=None, slots=None,
                ours_out=None, arms=("flashinfer", "flashinfer_q8", "triton", "fa4", "kv_write")):
    """phys_caches: list of HND physical [nb, 2, Hkv, bs, D].  Returns {arm: result}; `max_abs_vs_ours` added when ours_out given."""
    fp8 = phys_caches[0].dtype in (torch.uint8, torch.float8_e4m3fn)
    planes = [tuple(c.permute(0, 1, 3, 2, 4).unbind(1)) for c in phys_caches]
    Hkv = phys_caches[0].shape[2]
    max_len = int(lens.max().item())
    res = {}
    if "flashinfer" in arms:
        res["flashinfer_trtllm_gen_q_bf16"] = time_flashinfer(phys_caches, q, bt, lens, max_len, sm_scale, k_scale_t[0].item(),
                                                              v_scale_t[0].item(), dec_bytes, q_fp8=False)
    if "flashinfer_q8" in arms and fp8:
        res["flashinfer_trtllm_gen_q_fp8"] = time_flashinfer(phys_caches, q, bt, lens, max_len, sm_scale, k_scale_t[0].item(),
                                                             v_scale_t[0].item(), dec_bytes, q_fp8=True)
    if "triton" in arms:
        res["triton_unified_attention"] = time_triton(planes, q, bt, lens, max_len, sm_scale, k_scale_t, v_scale_t, dec_bytes, Hkv)
    if "fa4" in arms and not fp8:
        res["fa4_flash_attn_varlen"] = time_fa4(planes, q, bt, lens, max_len, sm_scale, dec_bytes)
    if "kv_write" in arms and kn is not None:
        res["vllm_reshape_and_cache_flash"] = time_vllm_kv_write(planes, kn, vn, slots, k_scale_t, v_scale_t, wr_bytes)
    if "flashinfer_q8" in arms and fp8:
        res["vllm_query_fp8_quant"] = time_vllm_query_quant(q)
    for k, v in res.items():
        if ours_out is not None:
            m = max_abs_vs(ours_out, v)
            if m is not None:
                v["max_abs_vs_ours"] = m
        elif isinstance(v, dict):
            v.pop("out", None)
    return res


def rel(x: float) -> float:
    return float(x) if math.isfinite(x) else float("nan")
