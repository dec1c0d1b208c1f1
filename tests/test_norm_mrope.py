"""q/k RMSNorm + M-RoPE + KV write in one launch (csrc/rope_kv_write.cu, torch.ops.b200attn.norm_rope_kv_write;
VERDICT r01 item 7 / SURVEY 8(f) rank 1: the Qwen3-Omni thinker's 3-axis positions and the Qwen3 q_norm / k_norm).

CPU: oracle.mrope_ref is pinned bit for bit to vLLM 0.22's MRotaryEmbedding.forward_native, oracle.rms_norm_ref to the native
rms_norm RMSNorm dispatches to (tests/golden/make_golden_mrope_norm.py -> vllm_mrope_rmsnorm_native.npz).
GPU: the kernel reproduces the M-RoPE goldens bit for bit (torch-native rounding) and equals the oracle bit for bit in the
compiled-op rounding; with the RMSNorm fused, q / k are within ONE unit in the last place of the reference (the fp32 sum of
squares is order dependent in the reference itself: block reduce on CUDA, pairwise on CPU) and the cache bytes within one e4m3
step where such a flip happened, identical elsewhere.
"""
import os

import numpy as np
import pytest
import torch

from oracle import oracle as O
from tests.test_rope import ACT, _bits, _dec, _enc, _t16

HERE = os.path.dirname(os.path.abspath(__file__))
MROPE = ["bf16_interleaved_24_20_20", "bf16_chunked_16_24_24", "fp16_interleaved_24_20_20"]


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(HERE, "golden", "vllm_mrope_rmsnorm_native.npz"))


def _meta(g, name):
    T, Hq, Hkv, D, rot, st, sh, sw, inter, max_pos = (int(x) for x in g[f"mrope_{name}_meta"])
    return T, Hq, Hkv, D, rot, (st, sh, sw), bool(inter), max_pos


@pytest.mark.parametrize("name", MROPE)
def test_oracle_mrope_bit_exact_vs_vllm_native(golden, name):
    g, act = golden, name[:4]
    T, Hq, Hkv, D, rot, sec, inter, _ = _meta(g, name)
    cache = _dec(g[f"mrope_{name}_cache"], act)
    q = _dec(g[f"mrope_{name}_q"], act).reshape(T, Hq, D)
    k = _dec(g[f"mrope_{name}_k"], act).reshape(T, Hkv, D)
    pos = g[f"mrope_{name}_pos"]
    q2 = O.mrope_ref(q, pos, cache, rot, True, act, sec, inter, each_op=True)
    k2 = O.mrope_ref(k, pos, cache, rot, True, act, sec, inter, each_op=True)
    assert np.array_equal(_enc(q2, act).reshape(T, -1), g[f"mrope_{name}_q_out"])
    assert np.array_equal(_enc(k2, act).reshape(T, -1), g[f"mrope_{name}_k_out"])


def test_mrope_axes_layouts():
    assert O.mrope_axes(16, (2, 3, 3), False).tolist() == [0, 0, 1, 1, 1, 2, 2, 2]
    assert O.mrope_axes(16, (2, 3, 3), True).tolist() == [0, 1, 2, 0, 1, 2, 0, 1][:8]
    ax = O.mrope_axes(128, (24, 20, 20), True)          # the thinker's layout: THWTHW... then the temporal tail
    assert ax[:6].tolist() == [0, 1, 2, 0, 1, 2] and ax[60:].tolist() == [0, 0, 0, 0]
    assert [(ax == a).sum() for a in range(3)] == [24, 20, 20]


@pytest.mark.parametrize("act", ["bf16", "fp16"])
def test_oracle_rms_norm_within_one_ulp_of_vllm_native(golden, act):
    g = golden
    x = _dec(g[f"rms_{act}_x"], act)
    w = _dec(g[f"rms_{act}_w"], act)
    ref = g[f"rms_{act}_y"]
    got = _enc(O.rms_norm_ref(x, w, float(g[f"rms_{act}_eps"]), act), act).reshape(ref.shape)
    diff = np.abs(got.astype(np.int64) - ref.astype(np.int64))      # same sign everywhere -> bit patterns are ordered
    assert diff.max() <= 1, int(diff.max())
    assert (diff != 0).mean() < 0.01      # in practice identical: only a last-bit tie of the fp32 statistics can differ


# ------------------------------------------------------------------------------------------------
# GPU
# ------------------------------------------------------------------------------------------------
def _empty(act):
    return torch.empty(0, dtype=ACT[act], device="cuda")


@pytest.mark.gpu
@pytest.mark.parametrize("name", MROPE)
def test_kernel_mrope_bit_exact_vs_vllm_golden(golden, name):
    import ht_vllm_omni_b200.ops  # noqa: F401
    g, act = golden, name[:4]
    T, Hq, Hkv, D, rot, sec, inter, _ = _meta(g, name)
    cache = _t16(g[f"mrope_{name}_cache"], act, (-1, rot))
    q = _t16(g[f"mrope_{name}_q"], act, (T, Hq, D))
    k = _t16(g[f"mrope_{name}_k"], act, (T, Hkv, D))
    v = torch.zeros_like(k)
    pos = torch.from_numpy(g[f"mrope_{name}_pos"]).cuda()
    bs, nb = 16, 4
    phys = torch.zeros((nb, 2, Hkv, bs, D), dtype=ACT[act], device="cuda")
    kc, vc = phys.permute(0, 1, 3, 2, 4).unbind(1)
    slots = torch.arange(T, dtype=torch.int64, device="cuda")
    one = torch.ones(1, dtype=torch.float32, device="cuda")
    torch.ops.b200attn.norm_rope_kv_write(q, k, v, pos, cache, True, _empty(act), _empty(act), 0.0, list(sec), inter, kc, vc, slots,
                                          one, one, True)   # torch-native rounding, as the golden's forward_native
    torch.cuda.synchronize()
    assert np.array_equal(_bits(q).reshape(T, -1), g[f"mrope_{name}_q_out"])
    assert np.array_equal(_bits(k).reshape(T, -1), g[f"mrope_{name}_k_out"])
    assert np.array_equal(_bits(kc.reshape(nb * bs, Hkv, D)[:T]).reshape(T, -1), g[f"mrope_{name}_k_out"])


def _problem(act, T=37, Hq=16, Hkv=2, D=128, seed=11):
    gen = torch.Generator().manual_seed(seed)
    qkv = (torch.randn(T, (Hq + 2 * Hkv) * D, generator=gen) * 2.0).to(ACT[act]).cuda()
    q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
    return q.view(T, Hq, D), k.view(T, Hkv, D), v.view(T, Hkv, D)


@pytest.mark.gpu
@pytest.mark.parametrize("act,kv,sec,inter,rot,each_op", [
    ("bf16", "fp8", (24, 20, 20), True, 128, False), ("bf16", "fp8", (16, 24, 24), False, 128, True),
    ("fp16", "fp16", (8, 12, 12), True, 64, False), ("bf16", "bf16", (24, 20, 20), True, 128, False)])
def test_mrope_kv_write_matches_oracle_bitwise(act, kv, sec, inter, rot, each_op):
    """M-RoPE without the norm: rotated q / k and the cache bytes equal the oracle bit for bit (both roundings)."""
    import ht_vllm_omni_b200.ops  # noqa: F401
    rng = np.random.default_rng(3)
    T, Hq, Hkv, D, bs, nb, max_pos = 37, 16, 2, 128, 16, 8, 5000
    fp8 = kv == "fp8"
    q, k, v = _problem(act)
    cache_np = O.rope_cos_sin_cache(rot, max_pos, 1e6, act)
    cache = torch.from_numpy(cache_np).to(ACT[act]).cuda()
    pos_np = rng.integers(0, max_pos, size=(3, T)).astype(np.int64)
    slots_np = rng.permutation(nb * bs)[:T].astype(np.int64)
    slots_np[[3, 20]] = -1
    ks, vs = np.array([0.02, 0.05], np.float32), np.array([0.04, 0.01], np.float32)
    phys = torch.zeros((nb, 2, Hkv, bs, D), dtype=torch.uint8 if fp8 else ACT[act], device="cuda")
    kc, vc = phys.permute(0, 1, 3, 2, 4).unbind(1)
    qn, kn, vn = q.float().cpu().numpy(), k.float().cpu().numpy(), v.float().cpu().numpy()
    q_ref = O.mrope_ref(qn, pos_np, cache_np, rot, True, act, sec, inter, each_op)
    k_ref = O.mrope_ref(kn, pos_np, cache_np, rot, True, act, sec, inter, each_op)
    kc_ref = np.zeros((nb, bs, Hkv, D), np.uint8 if fp8 else np.float32); vc_ref = np.zeros_like(kc_ref)
    O.kv_write_ref(k_ref, vn, kc_ref, vc_ref, slots_np, ks if fp8 else None, vs if fp8 else None, fp8)
    torch.ops.b200attn.norm_rope_kv_write(q, k, v, torch.from_numpy(pos_np).cuda(), cache, True, _empty(act), _empty(act), 0.0,
                                          list(sec), inter, kc, vc, torch.from_numpy(slots_np).cuda(),
                                          torch.from_numpy(ks).cuda(), torch.from_numpy(vs).cuda(), each_op)
    torch.cuda.synchronize()
    assert np.array_equal(q.float().cpu().numpy(), q_ref) and np.array_equal(k.float().cpu().numpy(), k_ref)
    got_k = kc.cpu().numpy() if fp8 else kc.float().cpu().numpy()
    got_v = vc.cpu().numpy() if fp8 else vc.float().cpu().numpy()
    assert np.array_equal(got_k, kc_ref) and np.array_equal(got_v, vc_ref)


@pytest.mark.gpu
@pytest.mark.parametrize("act,kv,mrope,rot", [("bf16", "fp8", True, 128), ("bf16", "fp8", False, 128), ("fp16", "fp16", False, 64),
                                              ("bf16", "bf16", True, 128)])
def test_norm_rope_kv_write_within_one_ulp_of_oracle(act, kv, mrope, rot):
    """RMSNorm fused in front of the rotation: every q / k element within one ulp of the oracle chain
    rms_norm_ref -> (m)rope_ref, V bytes exact, K bytes exact wherever the rotated key is."""
    import ht_vllm_omni_b200.ops  # noqa: F401
    rng = np.random.default_rng(9)
    T, Hq, Hkv, D, bs, nb, max_pos = 37, 16, 2, 128, 16, 8, 5000
    fp8 = kv == "fp8"
    sec, inter = (24, 20, 20), True
    q, k, v = _problem(act, seed=21)
    gen = torch.Generator().manual_seed(5)
    qw = (1.0 + 0.2 * torch.randn(D, generator=gen)).to(ACT[act]).cuda()
    kw = (1.0 + 0.2 * torch.randn(D, generator=gen)).to(ACT[act]).cuda()
    eps = 1e-6
    cache_np = O.rope_cos_sin_cache(rot, max_pos, 1e6, act)
    cache = torch.from_numpy(cache_np).to(ACT[act]).cuda()
    pos_np = rng.integers(0, max_pos, size=(3, T) if mrope else (T,)).astype(np.int64)
    slots_np = rng.permutation(nb * bs)[:T].astype(np.int64)
    slots_np[[5]] = -1
    ks, vs = np.array([0.02, 0.05], np.float32), np.array([0.04, 0.01], np.float32)
    phys = torch.zeros((nb, 2, Hkv, bs, D), dtype=torch.uint8 if fp8 else ACT[act], device="cuda")
    kc, vc = phys.permute(0, 1, 3, 2, 4).unbind(1)
    qn, kn, vn = q.float().cpu().numpy(), k.float().cpu().numpy(), v.float().cpu().numpy()
    rot_fn = (lambda x: O.mrope_ref(x, pos_np, cache_np, rot, True, act, sec, inter)) if mrope else \
             (lambda x: O.rope_ref(x, pos_np, cache_np, rot, True, act))
    q_ref = rot_fn(O.rms_norm_ref(qn, qw.float().cpu().numpy(), eps, act))
    k_ref = rot_fn(O.rms_norm_ref(kn, kw.float().cpu().numpy(), eps, act))
    torch.ops.b200attn.norm_rope_kv_write(q, k, v, torch.from_numpy(pos_np).cuda(), cache, True, qw, kw, eps, list(sec), inter, kc, vc,
                                          torch.from_numpy(slots_np).cuda(), torch.from_numpy(ks).cuda(), torch.from_numpy(vs).cuda(), False)
    torch.cuda.synchronize()
    mant = 7 if act == "bf16" else 10
    for got, ref in ((q.float().cpu().numpy(), q_ref), (k.float().cpu().numpy(), k_ref)):
        ulp = np.maximum(np.abs(ref), 2.0 ** -14) * 2.0 ** -mant
        assert np.all(np.abs(got - ref) <= ulp), float((np.abs(got - ref) / ulp).max())
        assert (got != ref).mean() < 0.02            # a normalised value landing on a rounding tie is rare
    assert np.array_equal(v.float().cpu().numpy(), vn)
    # the cache holds exactly kv_write's bytes of the key THE KERNEL produced (and of the untouched value)
    kc_ref = np.zeros((nb, bs, Hkv, D), np.uint8 if fp8 else np.float32); vc_ref = np.zeros_like(kc_ref)
    O.kv_write_ref(k.float().cpu().numpy(), vn, kc_ref, vc_ref, slots_np, ks if fp8 else None, vs if fp8 else None, fp8)
    got_k = kc.cpu().numpy() if fp8 else kc.float().cpu().numpy()
    got_v = vc.cpu().numpy() if fp8 else vc.float().cpu().numpy()
    assert np.array_equal(got_k, kc_ref) and np.array_equal(got_v, vc_ref)


@pytest.mark.gpu
def test_text_only_positions_through_the_mrope_entry_equal_plain_rope():
    """positions [3, T] with three equal rows (a text-only thinker batch) == b200attn_rope_kv_write with positions [T]."""
    import ht_vllm_omni_b200.ops  # noqa: F401
    T, Hq, Hkv, D, bs, nb = 37, 16, 2, 128, 16, 8
    q1, k1, v = _problem("bf16", seed=2)
    q2, k2 = q1.clone(), k1.clone()
    cache = torch.from_numpy(O.rope_cos_sin_cache(128, 4000, 1e6, "bf16")).to(torch.bfloat16).cuda()
    pos = torch.randint(0, 4000, (T,), generator=torch.Generator().manual_seed(1)).cuda()
    slots = torch.arange(T, dtype=torch.int64, device="cuda")
    s = torch.tensor([0.03], device="cuda")
    pa = torch.zeros((nb, 2, Hkv, bs, D), dtype=torch.uint8, device="cuda"); pb = torch.zeros_like(pa)
    ka, va = pa.permute(0, 1, 3, 2, 4).unbind(1); kb, vb = pb.permute(0, 1, 3, 2, 4).unbind(1)
    torch.ops.b200attn.rope_kv_write(q1, k1, v, pos, cache, True, ka, va, slots, s, s)
    e = torch.empty(0, dtype=torch.bfloat16, device="cuda")
    torch.ops.b200attn.norm_rope_kv_write(q2, k2, v, pos[None].repeat(3, 1).contiguous(), cache, True, e, e, 0.0, [24, 20, 20], True,
                                          kb, vb, slots, s, s)
    torch.cuda.synchronize()
    assert torch.equal(q1, q2) and torch.equal(k1, k2) and torch.equal(pa, pb)
