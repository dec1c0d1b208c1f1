"""Fused RoPE + KV write (csrc/rope_kv_write.cu, torch.ops.b200attn.rope_kv_write).

CPU: the oracle's rotation (oracle.rope_ref) is pinned bit-for-bit against vLLM 0.22's own RotaryEmbedding.forward_static
run on CPU (tests/golden/make_golden_rope.py -> tests/golden/vllm_rope_native.npz).
GPU: the CUDA kernel reproduces the golden q' / k' bit-for-bit, and on seeded problems the rotated q / k AND the cache bytes
equal oracle.rope_kv_write_ref (= rope_ref followed by the already-pinned kv_write_ref) bit-for-bit, for bf16 / fp16, neox /
GPT-J, partial rotary_dim, fp8 and 16-bit caches, strided qkv slices and padding slots.
"""
import os

import numpy as np
import pytest
import torch

from oracle import oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
CASES = ["bf16_neox", "bf16_gptj", "fp16_neox", "bf16_neox_rot64"]


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(HERE, "golden", "vllm_rope_native.npz"))


def _dec(u16, act):
    u16 = np.asarray(u16, dtype=np.uint16)
    return O.bf16_bits_to_f32(u16) if act == "bf16" else u16.view(np.float16).astype(np.float32)


def _enc(x, act):
    return O.f32_to_bf16_bits(x) if act == "bf16" else np.asarray(x, dtype=np.float32).astype(np.float16).view(np.uint16)


@pytest.mark.parametrize("name", CASES)
def test_oracle_rotation_bit_exact_vs_vllm_native(golden, name):
    g, act = golden, name[:4]
    T, Hq, Hkv, D, rot, neox, _ = (int(x) for x in g[name + "_meta"])
    cache = _dec(g[name + "_cache"], act)
    q = _dec(g[name + "_q"], act).reshape(T, Hq, D)
    k = _dec(g[name + "_k"], act).reshape(T, Hkv, D)
    q2 = O.rope_ref(q, g[name + "_pos"], cache, rot, bool(neox), act, each_op=True)
    k2 = O.rope_ref(k, g[name + "_pos"], cache, rot, bool(neox), act, each_op=True)
    assert np.array_equal(_enc(q2, act).reshape(T, -1), g[name + "_q_out"])
    assert np.array_equal(_enc(k2, act).reshape(T, -1), g[name + "_k_out"])


def test_oracle_cos_sin_cache_within_one_ulp_of_vllm(golden):
    # the table is an INPUT of the fused op; the helper only has to be a faithful generator for synthetic tests
    g = golden
    T, Hq, Hkv, D, rot, neox, max_pos = (int(x) for x in g["bf16_neox_meta"])
    ref = _dec(g["bf16_neox_cache"], "bf16")
    mine = O.rope_cos_sin_cache(rot, max_pos, float(g["base"]), "bf16")
    assert np.abs(ref - mine).max() <= 2.0 ** -8   # one bf16 ulp at magnitude 1


# ------------------------------------------------------------------------------------------------
# GPU
# ------------------------------------------------------------------------------------------------
ACT = {"bf16": torch.bfloat16, "fp16": torch.float16}


def _t16(u16, act, shape, device="cuda"):
    return torch.from_numpy(np.asarray(u16, dtype=np.uint16).view(np.int16).copy()).view(ACT[act]).reshape(shape).to(device)


def _bits(t):
    return t.detach().contiguous().view(torch.int16).cpu().numpy().view(np.uint16)


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_kernel_rotation_bit_exact_vs_vllm_golden(golden, name):
    import ht_vllm_omni_b200.ops  # noqa: F401  (registers torch.ops.b200attn)
    g, act = golden, name[:4]
    T, Hq, Hkv, D, rot, neox, _ = (int(x) for x in g[name + "_meta"])
    cache = _t16(g[name + "_cache"], act, (-1, rot))
    q = _t16(g[name + "_q"], act, (T, Hq, D))
    k = _t16(g[name + "_k"], act, (T, Hkv, D))
    v = torch.zeros_like(k)
    pos = torch.from_numpy(g[name + "_pos"]).cuda()
    bs, nb = 16, 4
    phys = torch.zeros((nb, 2, Hkv, bs, D), dtype=ACT[act], device="cuda")
    kc, vc = phys.permute(0, 1, 3, 2, 4).unbind(1)
    slots = torch.arange(T, dtype=torch.int64, device="cuda")
    one = torch.ones(1, dtype=torch.float32, device="cuda")
    torch.ops.b200attn.rope_kv_write(q, k, v, pos, cache, bool(neox), kc, vc, slots, one, one, True)   # torch-native rounding
    torch.cuda.synchronize()
    assert np.array_equal(_bits(q).reshape(T, -1), g[name + "_q_out"])
    assert np.array_equal(_bits(k).reshape(T, -1), g[name + "_k_out"])
    # 16-bit cache: the rotated key, bit for bit, at its slot
    got = kc.reshape(nb * bs, Hkv, D)[:T]
    assert np.array_equal(_bits(got).reshape(T, -1), g[name + "_k_out"])


@pytest.mark.gpu
@pytest.mark.parametrize("each_op", [False, True])
@pytest.mark.parametrize("act,kv,neox,rot,per_head", [
    ("bf16", "fp8", True, 128, False), ("bf16", "fp8", False, 128, True), ("fp16", "fp8", True, 64, True),
    ("bf16", "bf16", True, 128, False), ("fp16", "fp16", False, 32, False)])
def test_rope_kv_write_matches_oracle_bitwise(act, kv, neox, rot, per_head, each_op):
    import ht_vllm_omni_b200.ops  # noqa: F401
    rng = np.random.default_rng(5)
    T, Hq, Hkv, D, bs, nb, max_pos = 37, 16, 2, 128, 16, 8, 5000
    fp8 = kv == "fp8"
    gen = torch.Generator().manual_seed(11)
    # q / k / v are strided slices of one fused qkv activation, as vLLM hands them over
    qkv = torch.randn(T, (Hq + 2 * Hkv) * D, generator=gen).to(ACT[act]).cuda()
    q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
    q, k, v = q.view(T, Hq, D), k.view(T, Hkv, D), v.view(T, Hkv, D)
    cache_np = O.rope_cos_sin_cache(rot, max_pos, 1e6, act)
    cache = torch.from_numpy(cache_np).to(ACT[act]).cuda()
    pos_np = rng.integers(0, max_pos, size=T).astype(np.int64)
    slots_np = rng.permutation(nb * bs)[:T].astype(np.int64)
    slots_np[[3, 20]] = -1   # padding tokens: rotated, never written
    ks = (np.array([0.02, 0.05], np.float32) if per_head else np.array([0.03], np.float32))
    vs = (np.array([0.04, 0.01], np.float32) if per_head else np.array([0.02], np.float32))
    dt = torch.uint8 if fp8 else ACT[act]
    phys = torch.zeros((nb, 2, Hkv, bs, D), dtype=dt, device="cuda")
    kc, vc = phys.permute(0, 1, 3, 2, 4).unbind(1)
    # oracle on the same values
    qn, kn, vn = q.float().cpu().numpy(), k.float().cpu().numpy(), v.float().cpu().numpy()
    kc_ref = np.zeros((nb, bs, Hkv, D), np.uint8 if fp8 else np.float32)
    vc_ref = np.zeros_like(kc_ref)
    q_ref, k_ref = O.rope_kv_write_ref(qn, kn, vn, pos_np, cache_np, rot, neox, act, kc_ref, vc_ref, slots_np,
                                       ks if fp8 else None, vs if fp8 else None, fp8, each_op)
    torch.ops.b200attn.rope_kv_write(q, k, v, torch.from_numpy(pos_np).cuda(), cache, neox, kc, vc,
                                     torch.from_numpy(slots_np).cuda(), torch.from_numpy(ks).cuda(), torch.from_numpy(vs).cuda(), each_op)
    torch.cuda.synchronize()
    assert np.array_equal(q.float().cpu().numpy(), q_ref)
    assert np.array_equal(k.float().cpu().numpy(), k_ref)
    assert np.array_equal(v.float().cpu().numpy(), vn)   # value untouched
    if fp8:
        assert np.array_equal(kc.cpu().numpy(), kc_ref)
        assert np.array_equal(vc.cpu().numpy(), vc_ref)
    else:
        assert np.array_equal(kc.float().cpu().numpy(), kc_ref)
        assert np.array_equal(vc.float().cpu().numpy(), vc_ref)


@pytest.mark.gpu
def test_fused_equals_unfused_rotation_then_kv_write():
    """Idempotence of the split: rope_kv_write == (rope_kv_write into a scratch cache for the rotation) then kv_write."""
    import ht_vllm_omni_b200.ops  # noqa: F401
    T, Hq, Hkv, D, bs, nb = 64, 16, 2, 128, 16, 8
    gen = torch.Generator().manual_seed(3)
    q = torch.randn(T, Hq, D, generator=gen).bfloat16().cuda()
    k = torch.randn(T, Hkv, D, generator=gen).bfloat16().cuda()
    v = torch.randn(T, Hkv, D, generator=gen).bfloat16().cuda()
    cache = torch.from_numpy(O.rope_cos_sin_cache(D, 4096, 1e6, "bf16")).bfloat16().cuda()
    pos = torch.arange(T, dtype=torch.int64, device="cuda") * 7 % 4096
    slots = torch.randperm(nb * bs, generator=gen)[:T].to(torch.int64).cuda()
    sc = torch.full((1,), 0.02, device="cuda")
    a = torch.zeros((nb, 2, Hkv, bs, D), dtype=torch.uint8, device="cuda"); ka, va = a.permute(0, 1, 3, 2, 4).unbind(1)
    b = torch.zeros_like(a); kb, vb = b.permute(0, 1, 3, 2, 4).unbind(1)
    s = torch.zeros_like(a); ksc, vsc = s.permute(0, 1, 3, 2, 4).unbind(1)
    q1, k1 = q.clone(), k.clone()
    torch.ops.b200attn.rope_kv_write(q1, k1, v, pos, cache, True, ka, va, slots, sc, sc)
    q2, k2 = q.clone(), k.clone()
    torch.ops.b200attn.rope_kv_write(q2, k2, v, pos, cache, True, ksc, vsc, slots, sc, sc)   # rotation only matters here
    torch.ops.b200attn.kv_write(k2, v, kb, vb, slots, sc, sc)
    torch.cuda.synchronize()
    assert torch.equal(q1, q2) and torch.equal(k1, k2) and torch.equal(a, b)
