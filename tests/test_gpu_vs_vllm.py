"""GPU: the parity oracle OF RECORD -- the kernels the reference actually runs for this path, i.e. the
un-vendored vLLM dependency's own ops, executed on the same B200 on identical block tables and inputs:

  * torch.ops._C_cache_ops.reshape_and_cache_flash   (the compiled CUDA KV write, $VLLM/_custom_ops.py:2591-2610)
  * vllm.v1.attention.ops.triton_unified_attention.unified_attention   (2D and split-KV 3D paths)

They are used strictly as CHECKERS here (Triton is not part of our product path).  If the vLLM wheel cannot be
imported on the box the tests skip with the reason; nothing under /root/reference is needed at run time."""
import math

import numpy as np
import pytest
import torch

from tests.helpers import Case

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def vllm_ops():
    try:
        import vllm._custom_ops as cops  # noqa: F401  (loads vllm._C: needs the driver)
        from vllm.v1.attention.ops.triton_unified_attention import unified_attention
        from vllm.v1.kv_cache_interface import KVQuantMode
    except Exception as e:  # noqa: BLE001
        pytest.skip(f"vLLM kernels not importable on this box: {type(e).__name__}: {e}")
    if not hasattr(torch.ops, "_C_cache_ops") or not hasattr(torch.ops._C_cache_ops, "reshape_and_cache_flash"):
        pytest.skip("vllm._C_cache_ops.reshape_and_cache_flash missing from this wheel")
    return cops, unified_attention, KVQuantMode


@pytest.mark.parametrize("kv,dist,per_head", [("fp8", "normal", False), ("fp8", "heavy", False), ("bf16", "normal", False)])
def test_kv_write_bit_exact_vs_vllm_cuda_op(vllm_ops, kv, dist, per_head):
    cops, _, _ = vllm_ops
    import ht_vllm_omni_b200.ops  # noqa: F401

    c = Case([100, 33, 517, 1], 16, 2, kv=kv, dist=dist, per_head=per_head, layout="NHD").cuda()
    # specials: +-0, tiny, huge (saturation), inf
    c.d_key[0, 0, :8] = torch.tensor([0.0, -0.0, 1e-4, -1e-4, 3e4, -3e4, float("inf"), float("-inf")], dtype=torch.bfloat16)
    slots = c.d_slots.clone()
    slots[5] = -1
    torch.ops.b200attn.kv_write(c.d_key, c.d_value, c.kc, c.vc, slots, c.d_ks, c.d_vs)
    ref_cache = torch.zeros_like(c.cache)
    rk, rv = ref_cache.unbind(1)
    if kv == "fp8":
        cops.reshape_and_cache_flash(c.d_key, c.d_value, rk.view(torch.float8_e4m3fn), rv.view(torch.float8_e4m3fn), slots, "fp8",
                                     c.d_ks, c.d_vs)
    else:
        cops.reshape_and_cache_flash(c.d_key, c.d_value, rk, rv, slots, "auto", c.d_ks, c.d_vs)
    torch.cuda.synchronize()
    if kv == "fp8":
        assert torch.equal(c.kc.contiguous(), rk.contiguous()), int((c.kc != rk).sum())
        assert torch.equal(c.vc.contiguous(), rv.contiguous()), int((c.vc != rv).sum())
    else:
        assert torch.equal(c.kc.view(torch.int16).contiguous(), rk.view(torch.int16).contiguous())
        assert torch.equal(c.vc.view(torch.int16).contiguous(), rv.view(torch.int16).contiguous())


@pytest.mark.parametrize("kv,tol", [("fp8", 1e-2), ("bf16", 2e-3)])
@pytest.mark.parametrize("segs", [0, 16])
def test_decode_vs_vllm_triton_unified_attention(vllm_ops, kv, tol, segs):
    _, unified_attention, KVQuantMode = vllm_ops
    import ht_vllm_omni_b200.ops as ops

    lens = [1, 37, 100, 64, 513, 2048, 4096, 777]
    c = Case(lens, 16, 2, kv=kv, layout="NHD").cuda()
    torch.ops.b200attn.kv_write(c.d_key, c.d_value, c.kc, c.vc, c.d_slots, c.d_ks, c.d_vs)
    fp8 = kv == "fp8"
    B, Hq, Hkv, D = c.B, c.Hq, c.Hkv, c.D
    ref = torch.zeros(B, Hq, D, dtype=torch.bfloat16, device="cuda")
    cu = torch.arange(B + 1, dtype=torch.int32, device="cuda")
    kw = {}
    if segs:
        kw = dict(seq_threshold_3D=128, num_par_softmax_segments=segs,
                  softmax_segm_output=torch.zeros(B, Hq, segs, D, dtype=torch.float32, device="cuda"),
                  softmax_segm_max=torch.zeros(B, Hq, segs, dtype=torch.float32, device="cuda"),
                  softmax_segm_expsum=torch.zeros(B, Hq, segs, dtype=torch.float32, device="cuda"))
    kc = c.kc.view(torch.float8_e4m3fn) if fp8 else c.kc
    vc = c.vc.view(torch.float8_e4m3fn) if fp8 else c.vc
    unified_attention(c.d_q, kc, vc, ref, cu, 1, c.d_lens, int(max(lens)), 1.0 / math.sqrt(D), True, (-1, -1), c.d_bt, 0.0, None,
                      c.d_ks.expand(B, Hkv) if fp8 else None, c.d_vs.expand(B, Hkv) if fp8 else None,
                      kv_quant_mode=KVQuantMode.FP8_PER_TENSOR if fp8 else KVQuantMode.NONE, **kw)
    ws = ops.DecodeWorkspace("cuda", B, torch.uint8 if fp8 else torch.bfloat16)
    out = torch.zeros(B, Hq, D, dtype=torch.bfloat16, device="cuda")
    ws.make_plan(c.d_lens, Hq, Hkv, c.d_bt, c.bs)
    ws.decode(c.d_q, c.kc, c.vc, c.d_bt, c.d_lens, out, c.d_ks, c.d_vs, c.sm_scale)
    torch.cuda.synchronize()
    got, want = out.float().cpu().numpy(), ref.float().cpu().numpy()
    ulp = np.maximum(np.abs(want), 2.0 ** -14) * 2.0 ** -7     # both sides are rounded to bf16
    err = np.abs(got - want)
    assert np.all(err <= tol + ulp), f"max |ours - vLLM| = {err.max():.3e}"


@pytest.mark.parametrize("act,neox,kv", [("bf16", True, "fp8"), ("bf16", False, "bf16"), ("fp16", True, "fp8")])
def test_rope_kv_write_bit_exact_vs_vllm_cuda_ops(vllm_ops, act, neox, kv):
    """Fused launch vs the reference's two compiled ops: torch.ops._C.rotary_embedding ($VLLM/_custom_ops.py rotary_embedding,
    csrc/pos_encoding_kernels.cu) then reshape_and_cache_flash -- rotated q / k and the cache bytes, bit for bit."""
    cops, _, _ = vllm_ops
    if not hasattr(cops, "rotary_embedding"):
        pytest.skip("vllm._custom_ops.rotary_embedding missing from this wheel")
    import ht_vllm_omni_b200.ops  # noqa: F401
    from oracle import oracle as O

    dt = {"bf16": torch.bfloat16, "fp16": torch.float16}[act]
    T, Hq, Hkv, D, bs, nb, max_pos = 200, 16, 2, 128, 16, 16, 8192
    gen = torch.Generator().manual_seed(21)
    q = torch.randn(T, Hq * D, generator=gen).to(dt).cuda()
    k = torch.randn(T, Hkv * D, generator=gen).to(dt).cuda()
    v = torch.randn(T, Hkv * D, generator=gen).to(dt).cuda()
    cache = torch.from_numpy(O.rope_cos_sin_cache(D, max_pos, 1e6, act)).to(dt).cuda()
    pos = torch.randint(0, max_pos, (T,), generator=gen, dtype=torch.int64).cuda()
    slots = torch.randperm(nb * bs, generator=gen)[:T].to(torch.int64).cuda()
    slots[9] = -1
    ks = torch.full((1,), 0.03, device="cuda"); vs = torch.full((1,), 0.02, device="cuda")
    fp8 = kv == "fp8"
    mine = torch.zeros((nb, 2, bs, Hkv, D), dtype=torch.uint8 if fp8 else dt, device="cuda")
    ref = torch.zeros_like(mine)
    mk, mv = mine.unbind(1)
    rk, rv = ref.unbind(1)
    q1, k1 = q.clone(), k.clone()
    torch.ops.b200attn.rope_kv_write(q1.view(T, Hq, D), k1.view(T, Hkv, D), v.view(T, Hkv, D), pos, cache, neox, mk, mv, slots, ks, vs)
    q2, k2 = q.clone(), k.clone()
    cops.rotary_embedding(pos, q2, k2, D, cache, neox)
    if fp8:
        cops.reshape_and_cache_flash(k2.view(T, Hkv, D), v.view(T, Hkv, D), rk.view(torch.float8_e4m3fn), rv.view(torch.float8_e4m3fn),
                                     slots, "fp8", ks, vs)
    else:
        cops.reshape_and_cache_flash(k2.view(T, Hkv, D), v.view(T, Hkv, D), rk, rv, slots, "auto", ks, vs)
    torch.cuda.synchronize()
    assert torch.equal(q1.view(torch.int16), q2.view(torch.int16)), int((q1.view(torch.int16) != q2.view(torch.int16)).sum())
    assert torch.equal(k1.view(torch.int16), k2.view(torch.int16))
    if fp8:
        assert torch.equal(mine, ref), int((mine != ref).sum())
    else:
        assert torch.equal(mine.view(torch.int16), ref.view(torch.int16))
