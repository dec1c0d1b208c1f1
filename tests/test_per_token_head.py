"""kv_cache_dtype "fp8_per_token_head" (SURVEY 8(f) rank 4): one dynamic FP8 scale per (token, head) in float32 side caches.

CPU   oracle.kv_write_per_token_head_ref == vLLM's _reshape_cache_per_token_head run under the Triton interpreter
      (tests/golden/make_golden_per_token_head.py): scales bit for bit, bytes bit for bit except where the interpreter's
      own fp8 cast loses a rounding carry (the compiled kernel does not; see _equal_up_to_interpreter_carry).
GPU   b200attn_kv_write_per_token_head (IEEE-division flag) == that golden and == the oracle on seeded bf16 problems, bit for bit;
      its default arithmetic (div.full.f32) == vLLM's compiled Triton kernel on the box, bit for bit;
      b200attn_decode_per_token_head == the fp64 definition on the dequantised cache within north_star's FP8 tolerance, through
      the backend's own block layout (ops.per_token_head_views), cut rows, ragged batches, 16 heads per kv head.
"""
import os

import numpy as np
import pytest
import torch

from oracle import oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(HERE, "golden", "vllm_triton_interp_per_token_head.npz"))


def _equal_up_to_interpreter_carry(mine, interp):
    """The Triton INTERPRETER's float32 -> e4m3 cast is not the hardware's: it drops the carry when rounding reaches the next
    power of two (-125.95 -> -64 instead of -128), rounds ties away from zero (272 -> 288, not the even 256) and mis-rounds
    subnormals -- under 1 % of these elements.  The compiled kernel (pinned bit for bit on the box below) and IEEE
    round-to-nearest-even agree with each other.  So: bytes identical everywhere else, and every mismatch is one of those
    three shapes (same sign; power of two of twice the magnitude, or the neighbouring code, or a subnormal within two codes)."""
    diff = mine != interp
    if not diff.any():
        return True
    a, b = mine[diff].astype(np.int64), interp[diff].astype(np.int64)
    same_sign = (a & 0x80) == (b & 0x80)
    carry = ((a & 7) == 0) & (O.e4m3_decode(a.astype(np.uint8)) == 2.0 * O.e4m3_decode(b.astype(np.uint8)))
    tie = np.abs(a - b) == 1
    sub = ((a & 0x7f) < 8) & (np.abs(a - b) <= 2)
    return bool((same_sign & (carry | tie | sub)).all()) and diff.mean() < 0.02


def test_oracle_write_bit_exact_vs_vllm_triton_interpreter(golden):
    g = golden
    T, Hkv, D, bs, nb = (int(x) for x in g["meta"])
    kc = np.zeros((nb, bs, Hkv, D), np.uint8); vc = np.zeros_like(kc)
    ksc = np.ones((nb, bs, Hkv), np.float32); vsc = np.ones_like(ksc)
    O.kv_write_per_token_head_ref(g["key"].astype(np.float32), g["value"].astype(np.float32), kc, vc, ksc, vsc, g["slots"])
    assert np.array_equal(ksc, g["k_scale_cache"]) and np.array_equal(vsc, g["v_scale_cache"])
    assert _equal_up_to_interpreter_carry(kc, g["k_cache"]) and _equal_up_to_interpreter_carry(vc, g["v_cache"])
    assert g["k_scale_cache"].min() == np.float32(1e-6)          # the all-zero head hit the floor
    assert np.isclose(g["v_scale_cache"], 300.0 / 448.0).any()   # the outlier set its head's scale


def _alloc(nb, bs, Hkv, D):
    """A cache as vLLM hands it to this backend in per-token-head mode: [nb, 2, bs, Hkv, D + 4] bytes, HND stride order."""
    import ht_vllm_omni_b200.ops as ops

    raw = torch.zeros(nb * 2 * bs * Hkv * (D + 4), dtype=torch.uint8, device="cuda")
    kv_cache = raw.view(nb, 2, Hkv, bs, D + 4).permute(0, 1, 3, 2, 4)
    return kv_cache, ops.per_token_head_views(kv_cache, Hkv, D)


@pytest.mark.gpu
def test_views_tile_the_block_exactly():
    kv_cache, (k, v, ks, vs) = _alloc(5, 32, 2, 128)
    k.fill_(1); v.fill_(2); ks.fill_(3.0); vs.fill_(4.0)
    flat = kv_cache.permute(0, 1, 3, 2, 4).reshape(5, -1)      # physical order: one row per block
    blk = 2 * 32 * 2 * 132
    assert flat.shape[1] == blk
    n = 2 * 32 * 128
    assert bool((flat[:, :n] == 1).all()) and bool((flat[:, n:2 * n] == 2).all())
    f = flat[:, 2 * n:].contiguous().view(torch.float32)
    assert bool((f[:, :64] == 3.0).all()) and bool((f[:, 64:] == 4.0).all())
    assert ks.stride(1) == 1 and k.stride() == v.stride() and k.stride(1) == 128


@pytest.mark.gpu
def test_kernel_write_ieee_bit_exact_vs_golden_and_oracle(golden):
    import ht_vllm_omni_b200.ops  # noqa: F401
    g = golden
    T, Hkv, D, bs, nb = (int(x) for x in g["meta"])
    _, (k, v, ks, vs) = _alloc(nb, bs, Hkv, D)
    ks.fill_(1.0); vs.fill_(1.0)
    torch.ops.b200attn.kv_write_per_token_head(torch.from_numpy(g["key"]).cuda(), torch.from_numpy(g["value"]).cuda(), k, v, ks, vs,
                                               torch.from_numpy(g["slots"]).cuda(), True)
    torch.cuda.synchronize()
    assert np.array_equal(ks.cpu().numpy(), g["k_scale_cache"]) and np.array_equal(vs.cpu().numpy(), g["v_scale_cache"])
    assert _equal_up_to_interpreter_carry(k.cpu().numpy(), g["k_cache"]) and _equal_up_to_interpreter_carry(v.cpu().numpy(), g["v_cache"])
    # seeded bf16 problem, strided key / value slices of a fused qkv activation, padding slots
    gen = torch.Generator().manual_seed(3)
    T, Hq, Hkv, bs, nb = 70, 16, 2, 16, 6
    qkv = (torch.randn(T, (Hq + 2 * Hkv) * D, generator=gen) * 1.3).to(torch.bfloat16).cuda()
    key = qkv[:, Hq * D:(Hq + Hkv) * D].view(T, Hkv, D); value = qkv[:, (Hq + Hkv) * D:].view(T, Hkv, D)
    slots = torch.randperm(nb * bs, generator=gen)[:T].to(torch.int64)
    slots[[2, 50]] = -1
    _, (k, v, ks, vs) = _alloc(nb, bs, Hkv, D)
    torch.ops.b200attn.kv_write_per_token_head(key, value, k, v, ks, vs, slots.cuda(), True)
    torch.cuda.synchronize()
    kc = np.zeros((nb, bs, Hkv, D), np.uint8); vc = np.zeros_like(kc)
    ksc = np.zeros((nb, bs, Hkv), np.float32); vsc = np.zeros_like(ksc)
    O.kv_write_per_token_head_ref(key.float().cpu().numpy(), value.float().cpu().numpy(), kc, vc, ksc, vsc, slots.numpy())
    assert np.array_equal(ks.cpu().numpy(), ksc) and np.array_equal(vs.cpu().numpy(), vsc)
    assert np.array_equal(k.cpu().numpy(), kc) and np.array_equal(v.cpu().numpy(), vc)


@pytest.mark.gpu
def test_kernel_write_default_bit_exact_vs_vllm_compiled_triton_kernel():
    """The reference of record for this row exists only as a Triton kernel; compiled, its float32 `/` is div.full.f32."""
    import ht_vllm_omni_b200.ops  # noqa: F401
    try:
        from vllm.v1.attention.ops.triton_reshape_and_cache_flash import triton_reshape_and_cache_flash_per_token_head_quant as ref
    except Exception as e:  # noqa: BLE001
        pytest.skip(f"vLLM's per-token-head kernel is not importable on this box: {e}")
    gen = torch.Generator().manual_seed(8)
    T, Hkv, D, bs, nb = 1000, 4, 128, 16, 70
    key = (torch.randn(T, Hkv, D, generator=gen) * 2.1).to(torch.bfloat16).cuda()
    value = (torch.randn(T, Hkv, D, generator=gen) * 0.4).to(torch.bfloat16).cuda()
    key[17, 2] = 0
    slots = torch.randperm(nb * bs, generator=gen)[:T].to(torch.int64).cuda()
    slots[5] = -1
    rk = torch.zeros(nb, bs, Hkv, D, dtype=torch.float8_e4m3fn, device="cuda"); rv = torch.zeros_like(rk)
    rks = torch.ones(nb, bs, Hkv, dtype=torch.float32, device="cuda"); rvs = torch.ones_like(rks)
    ref(key, value, rk, rv, rks, rvs, slots)
    _, (k, v, ks, vs) = _alloc(nb, bs, Hkv, D)
    ks.fill_(1.0); vs.fill_(1.0)
    torch.ops.b200attn.kv_write_per_token_head(key, value, k, v, ks, vs, slots)
    torch.cuda.synchronize()
    assert torch.equal(ks, rks) and torch.equal(vs, rvs), "scales differ from vLLM's compiled kernel"
    assert torch.equal(k, rk.view(torch.uint8)) and torch.equal(v, rv.view(torch.uint8)), "cache bytes differ from vLLM's compiled kernel"


CASES = [("ragged", [1, 37, 100, 0, 64, 513, 2048], 16, 2, None), ("cut_rows", [300, 700, 1500, 90], 16, 2, 7),
         ("thinker_heads", [777, 1500, 64], 32, 4, None), ("G16_two_accumulator_sets", [300, 700], 16, 1, 3)]


@pytest.mark.gpu
@pytest.mark.parametrize("name,lens,Hq,Hkv,ctas", CASES, ids=[c[0] for c in CASES])
def test_decode_matches_fp64_definition_on_dequantised_cache(name, lens, Hq, Hkv, ctas):
    import ht_vllm_omni_b200.ops as ops
    from tests.test_gpu_parity import assert_close

    D, bs = 128, 16
    lens = np.asarray(lens, np.int32)
    B = len(lens)
    bt_np, nb = O.make_block_table(lens, bs, seed=13)
    gen = torch.Generator().manual_seed(21)
    Tt = int(lens.sum())
    # heads and tokens of very different magnitude: what per-token scales are for
    mag = torch.exp(torch.randn(Tt, Hkv, 1, generator=gen) * 1.5)
    key = (torch.randn(Tt, Hkv, D, generator=gen) * mag).to(torch.bfloat16).cuda()
    value = (torch.randn(Tt, Hkv, D, generator=gen) * mag.flip(0)).to(torch.bfloat16).cuda()
    q = torch.randn(B, Hq, D, generator=gen).to(torch.bfloat16).cuda()
    pos = [(r, p) for r in range(B) for p in range(int(lens[r]))]
    slots = torch.from_numpy(O.slots_for_positions(bt_np, pos, bs)).cuda()
    _, (k, v, ks, vs) = _alloc(nb, bs, Hkv, D)
    # stale scales past seq_len must never reach the output: poison everything first
    ks.fill_(float("nan")); vs.fill_(float("inf"))
    torch.ops.b200attn.kv_write_per_token_head(key, value, k, v, ks, vs, slots)
    ws = ops.DecodeWorkspace("cuda", B, torch.uint8, num_ctas=ctas, group_size=Hq // Hkv)
    d_lens, d_bt = torch.from_numpy(lens).cuda(), torch.from_numpy(bt_np).cuda()
    ws.make_plan(d_lens, Hq, Hkv, d_bt, bs)
    out = torch.full((B, Hq, D), float("nan"), dtype=torch.bfloat16, device="cuda")
    ws.decode_per_token_head(q, k, v, ks, vs, d_bt, d_lens, out, D ** -0.5)
    torch.cuda.synchronize()
    ks_np, vs_np = ks.cpu().numpy(), vs.cpu().numpy()
    kf = O.dequant_per_token_head(k.cpu().numpy(), np.nan_to_num(ks_np, nan=0.0, posinf=0.0))
    vf = O.dequant_per_token_head(v.cpu().numpy(), np.nan_to_num(vs_np, nan=0.0, posinf=0.0))
    ref = O.paged_decode_exact(q.float().cpu().numpy(), kf, vf, bt_np, lens, D ** -0.5, fp8=False)
    live = lens > 0
    got = out.float().cpu().numpy()
    vmax = float(np.abs(vf).max())
    assert_close(got, ref.astype(np.float32), "fp8", "bf16", live, vmax=vmax if vmax > 6 else None, exact=True, tag=f"per_token_head_{name}")


@pytest.mark.gpu
@pytest.mark.parametrize("q_lens", [[1, 1, 1, 1], [1, 3, 2, 7]], ids=["pure_decode", "mixed_multi_token_rows"])
def test_backend_step_with_per_token_head_cache(q_lens):
    """The plugin classes with kv_cache_dtype="fp8_per_token_head": vLLM-style allocation of the padded shape, the step's
    tokens written by forward() (forward_includes_kv_cache_update), every query token a decode row; the attention equals the
    fp64 definition on what the cache then holds."""
    import types

    import vllm.v1.attention.backend as vb

    from ht_vllm_omni_b200 import backend as BK
    from tests.test_gpu_backend import _alloc_like_vllm, _cfg

    be = BK.B200PagedAttentionBackend
    Hq, Hkv, D, bs = 16, 2, 128, 16
    assert be.supports_kv_cache_dtype(BK.PER_TOKEN_HEAD) and be.get_kv_cache_shape(3, bs, Hkv, D, BK.PER_TOKEN_HEAD) == (3, 2, bs, Hkv, D + 4)
    ctx = np.array([300, 40, 0, 1000], np.int32)
    q_lens = np.array(q_lens, np.int32)
    seq = ctx + q_lens
    B_, T = len(seq), int(q_lens.sum())
    bt_np, nb = O.make_block_table(seq, bs, seed=5)
    gen = torch.Generator().manual_seed(17)
    rnd = lambda *s: (torch.randn(*s, generator=gen) * torch.exp(torch.randn(s[0], s[1], 1, generator=gen))).to(torch.bfloat16)  # noqa: E731
    hist_k = [rnd(int(L), Hkv, D) for L in seq]; hist_v = [rnd(int(L), Hkv, D) for L in seq]
    kv_cache = _alloc_like_vllm(be, nb, bs, Hkv, D, BK.PER_TOKEN_HEAD, torch.uint8)
    assert tuple(kv_cache.shape) == (nb, 2, bs, Hkv, D + 4)
    layer = types.SimpleNamespace(_k_scale=torch.ones(1, device="cuda"), _v_scale=torch.ones(1, device="cuda"))
    impl = be.get_impl_cls()(Hq, D, D ** -0.5, Hkv, None, None, BK.PER_TOKEN_HEAD)
    pos_ctx = [(r, p) for r in range(B_) for p in range(int(ctx[r]))]
    ck = torch.cat([hist_k[r][: ctx[r]] for r in range(B_)]).reshape(-1, Hkv * D).cuda()
    cv = torch.cat([hist_v[r][: ctx[r]] for r in range(B_)]).reshape(-1, Hkv * D).cuda()
    impl.do_kv_cache_update(layer, ck, cv, kv_cache, torch.from_numpy(O.slots_for_positions(bt_np, pos_ctx, bs)).cuda())
    new_k = torch.cat([hist_k[r][ctx[r]:] for r in range(B_)]).reshape(T, -1).cuda()
    new_v = torch.cat([hist_v[r][ctx[r]:] for r in range(B_)]).reshape(T, -1).cuda()
    q = torch.randn(T, Hq * D, generator=gen).to(torch.bfloat16).cuda()
    pos_new = [(r, p) for r in range(B_) for p in range(int(ctx[r]), int(seq[r]))]
    slots = torch.from_numpy(O.slots_for_positions(bt_np, pos_new, bs)).cuda()
    qsl = np.concatenate([[0], np.cumsum(q_lens)]).astype(np.int32)
    cm = vb.CommonAttentionMetadata(
        query_start_loc=torch.from_numpy(qsl).cuda(), query_start_loc_cpu=torch.from_numpy(qsl), seq_lens=torch.from_numpy(seq).cuda(),
        num_reqs=B_, num_actual_tokens=T, max_query_len=int(q_lens.max()), max_seq_len=int(seq.max()),
        block_table_tensor=torch.from_numpy(bt_np).cuda(), slot_mapping=slots)
    spec = types.SimpleNamespace(block_size=bs, dtype=torch.uint8, kv_quant_mode=types.SimpleNamespace(is_per_token_head=True))
    builder = be.get_builder_cls()(spec, ["layer.0"], _cfg(Hq, Hkv, D, 64), torch.device("cuda"))
    assert builder.per_token_head and not builder.use_prefill_kernel
    md = builder.build(0, cm)
    if not be.forward_includes_kv_cache_update:
        impl.do_kv_cache_update(layer, new_k, new_v, kv_cache, slots)
    output = torch.zeros(T, Hq * D, dtype=torch.bfloat16, device="cuda")
    impl.forward(layer, q, new_k, new_v, kv_cache, md, output)
    torch.cuda.synchronize()
    import ht_vllm_omni_b200.ops as ops

    k, v, ks, vs = ops.per_token_head_views(kv_cache, Hkv, D)
    # every position of every sequence was written (context by the hook, this step by forward): scales are finite and positive there
    all_slots = O.slots_for_positions(bt_np, [(r, p) for r in range(B_) for p in range(int(seq[r]))], bs)
    ks_np, vs_np = ks.cpu().numpy(), vs.cpu().numpy()
    assert (ks_np.reshape(-1, Hkv)[all_slots] > 0).all() and (vs_np.reshape(-1, Hkv)[all_slots] > 0).all()
    kf = O.dequant_per_token_head(k.cpu().numpy(), ks_np); vf = O.dequant_per_token_head(v.cpu().numpy(), vs_np)
    row_req = np.repeat(np.arange(B_), q_lens)
    row_len = np.concatenate([ctx[r] + 1 + np.arange(q_lens[r]) for r in range(B_)]).astype(np.int32)
    ref = O.paged_decode_exact(q.view(T, Hq, D).float().cpu().numpy(), kf, vf, bt_np[row_req], row_len, D ** -0.5, fp8=False)
    got = output.view(T, Hq, D).float().cpu().numpy()
    err = np.abs(got - ref)
    vmax = float(np.abs(vf).max())
    tol = 1e-2 * max(1.0, vmax / 6.0)
    assert np.all(err <= tol + 0.5 * np.maximum(np.abs(ref), 2.0 ** -14) * 2.0 ** -7), float(err.max())
