"""GPU parity of the prefill / mixed-batch kernel (b200attn_prefill, q_len > 1): the tcgen05 path through
torch.ops.b200attn.paged_prefill -> C ABI against the CPU oracle (oracle.paged_prefill_ref: reference-style rounding points,
and its float64 `exact` form) on the same seeded inputs, and against vLLM's own Triton unified_attention on the box.

Tolerance: BASELINE.json north_star, FP8 KV: max-abs 1e-2 (+ one ulp of the 16-bit output at the reference's magnitude)."""
import numpy as np
import pytest
import torch

from oracle import oracle as O
from tests.helpers import ACT, Case, to_np_f32

pytestmark = pytest.mark.gpu
TOL = 1e-2


def close(got, ref, act, what):
    assert not np.isnan(got).any(), f"{what}: NaN in CUDA output"
    ulp = np.maximum(np.abs(ref), 2.0 ** -14) * 2.0 ** -(7 if act == "bf16" else 10)
    err = np.abs(got - ref)
    bad = err > TOL + ulp
    assert not bad.any(), f"{what}: max err {err.max():.3e} (tol {TOL} + 1 ulp), {int(bad.sum())} elements out"
    return float(err.max())


@pytest.fixture(scope="module")
def ops():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import ht_vllm_omni_b200.ops as ops

    ops.load_library()
    return ops


class PrefillCase:
    """`ctx[r]` cached tokens + a chunk of `qlen[r]` new ones per request (both already written to the cache, as
    do_kv_cache_update precedes forward)."""

    def __init__(self, ctx, qlen, Hq, Hkv, act="bf16", seed=0, per_head=False, bs=16):
        self.ctx, self.qlen = np.asarray(ctx, np.int32), np.asarray(qlen, np.int32)
        self.c = Case(list(self.ctx + self.qlen), Hq, Hkv, kv="fp8", act=act, seed=seed, per_head=per_head, bs=bs).cuda()
        c = self.c
        g = torch.Generator().manual_seed(777 + seed)
        self.T = int(self.qlen.sum())
        self.q = torch.randn(self.T, Hq, c.D, generator=g).to(ACT[act])
        self.qsl = np.concatenate([[0], np.cumsum(self.qlen)]).astype(np.int32)
        torch.ops.b200attn.kv_write(c.d_key, c.d_value, c.kc, c.vc, c.d_slots, c.d_ks, c.d_vs)
        torch.cuda.synchronize()
        self.kc_np, self.vc_np = c.cache_to_np(c.kc), c.cache_to_np(c.vc)

    def run(self, ops, qsl=None, first_token=0, q=None, out=None):
        c = self.c
        d_q = self.q.cuda() if q is None else q
        d_qsl = torch.from_numpy(self.qsl if qsl is None else qsl).cuda()
        tp = torch.zeros(len(self.qlen) + 1, dtype=torch.int32, device="cuda")
        out = torch.full_like(d_q, float("nan")) if out is None else out
        torch.ops.b200attn.prefill_plan(d_qsl, tp, c.Hq, c.Hkv)
        torch.ops.b200attn.paged_prefill(d_q, c.kc, c.vc, c.d_bt, c.d_lens, d_qsl, tp, out, c.d_ks, c.d_vs, c.sm_scale, first_token)
        torch.cuda.synchronize()
        return out, tp

    def oracle(self, exact=False, rows=None):
        c = self.c
        return O.paged_prefill_ref(to_np_f32(self.q), self.qsl, self.kc_np, self.vc_np, c.bt_np, c.lens, c.sm_scale,
                                   c.k_scale.numpy(), c.v_scale.numpy(), fp8=True, act=c.act, exact=exact, rows=rows)


CASES = [
    # name, ctx, qlen, Hq, Hkv, act, per_head
    ("one_tile_no_context", [0], [16], 8, 1, "bf16", False),
    ("one_token_chunks_are_decode", [5, 300], [1, 1], 8, 1, "bf16", False),
    ("partial_tile", [0], [5], 8, 1, "bf16", False),
    ("two_kv_tiles", [0], [200], 8, 1, "bf16", False),
    ("context_plus_chunk", [300], [40], 8, 1, "bf16", False),
    ("ragged_mixed", [0, 17, 500, 1000, 3], [33, 1, 128, 7, 250], 16, 2, "bf16", True),
    ("g4", [100, 0], [70, 129], 8, 2, "bf16", False),
    ("g1_mha", [100, 0], [130, 260], 2, 2, "bf16", False),
    ("g16", [64, 0], [9, 40], 16, 1, "bf16", True),
    ("g7_qwen2", [50, 0], [37, 19], 14, 2, "bf16", False),
    ("g32_split_group", [40], [21], 32, 1, "bf16", False),
    ("fp16_activations", [200, 0], [64, 100], 16, 2, "fp16", True),
    ("chunk_of_long_context", [4096], [512], 8, 1, "bf16", False),
    ("block_size_64", [100, 0], [70, 129], 8, 2, "bf16", False),
]


@pytest.mark.parametrize("name,ctx,qlen,Hq,Hkv,act,per_head", CASES, ids=[c[0] for c in CASES])
def test_prefill_matches_the_oracle(ops, name, ctx, qlen, Hq, Hkv, act, per_head):
    pc = PrefillCase(ctx, qlen, Hq, Hkv, act=act, per_head=per_head, bs=64 if name == "block_size_64" else 16)
    out, tp = pc.run(ops)
    tpt = ops.prefill_tile_tokens(Hq, Hkv)
    assert tp.cpu().tolist() == np.concatenate([[0], np.cumsum(-(-pc.qlen // tpt))]).tolist()
    got = to_np_f32(out)
    close(got, pc.oracle(), act, f"{name} vs reference-style oracle")
    close(got, pc.oracle(exact=True), act, f"{name} vs float64 definition")


def test_one_token_chunks_equal_the_decode_kernel(ops):
    """q_len = 1 rows are plain decode: both kernels must agree with each other (and the oracle) on them."""
    pc = PrefillCase([37, 999, 4000], [1, 1, 1], 16, 2, per_head=True)
    out, _ = pc.run(ops)
    c = pc.c
    ws = ops.DecodeWorkspace("cuda", c.B, torch.uint8)
    ws.make_plan(c.d_lens, c.Hq, c.Hkv, c.d_bt, c.bs)
    dec = torch.empty_like(out)
    ws.decode(pc.q.cuda(), c.kc, c.vc, c.d_bt, c.d_lens, dec, c.d_ks, c.d_vs, c.sm_scale)
    torch.cuda.synchronize()
    assert float((dec.float() - out.float()).abs().max()) <= 2 * 2.0 ** -7 * float(dec.float().abs().max())


def test_requests_without_query_tokens_and_pre_offset_rows(ops):
    """A mixed batch where decode rows come first: the prefill op gets the rows of requests [2:), the full q / out
    tensors and first_token = 0; rows of other requests must not be touched.  Also a request with q_len = 0."""
    pc = PrefillCase([10, 20, 0, 64, 5], [1, 1, 50, 0, 31], 8, 1)
    c = pc.c
    d_q = pc.q.cuda()
    out = torch.full_like(d_q, 7.0)
    d_qsl = torch.from_numpy(pc.qsl).cuda()
    tp = torch.zeros(4, dtype=torch.int32, device="cuda")
    torch.ops.b200attn.prefill_plan(d_qsl[2:], tp, c.Hq, c.Hkv)
    torch.ops.b200attn.paged_prefill(d_q, c.kc, c.vc, c.d_bt[2:], c.d_lens[2:], d_qsl[2:], tp, out, c.d_ks, c.d_vs, c.sm_scale, 0)
    torch.cuda.synchronize()
    got = to_np_f32(out)
    assert (got[:2] == 7.0).all(), "decode rows were overwritten by the prefill kernel"
    close(got[2:], pc.oracle()[2:], "bf16", "rows of the prefill requests")


def test_stale_cache_bytes_past_seq_len_do_not_leak(ops):
    """Slots past seq_len inside the last block, and whole blocks past it, hold NaN patterns (0x7f): they must not reach
    the output through 0 * NaN in the second GEMM."""
    pc = PrefillCase([0, 40], [37, 9], 8, 1)
    c = pc.c
    kc_o, vc_o = pc.kc_np.copy(), pc.vc_np.copy()
    c.kc.fill_(0x7F); c.vc.fill_(0x7F)
    torch.ops.b200attn.kv_write(c.d_key, c.d_value, c.kc, c.vc, c.d_slots, c.d_ks, c.d_vs)
    out, _ = pc.run(ops)
    close(to_np_f32(out), pc.oracle(), "bf16", "poisoned cache")


def test_full_size_chunked_prefill_sampled_rows(ops):
    """BASELINE-sized chunk (2048 new tokens on an 30k-token context, thinker heads 32 / 4): every output row is finite and a
    seeded sample of rows matches the oracle."""
    pc = PrefillCase([30000], [2048], 32, 4, per_head=True)
    out, _ = pc.run(ops)
    got = to_np_f32(out)
    assert np.isfinite(got).all()
    rows = [0, 1, 127, 128, 1000, 2047]
    ref = pc.oracle(rows=rows)
    close(got[rows], ref[rows], "bf16", "sampled rows of the full-size chunk")


def test_matches_vllm_triton_unified_attention(ops):
    """The reference's own prefill kernel (vLLM Triton unified_attention) on the same cache, run on this box."""
    ua = pytest.importorskip("vllm.v1.attention.ops.triton_unified_attention")
    pc = PrefillCase([0, 17, 500, 1000, 3], [33, 1, 128, 7, 250], 16, 2, per_head=False, seed=3)
    c = pc.c
    out, _ = pc.run(ops)
    d_q = pc.q.cuda()
    ref = torch.empty_like(d_q)
    kview, vview = c.kc.view(torch.float8_e4m3fn), c.vc.view(torch.float8_e4m3fn)
    try:
        from vllm.v1.kv_cache_interface import KVQuantMode

        R = len(pc.qlen)
        ua.unified_attention(d_q, kview, vview, ref, torch.from_numpy(pc.qsl).cuda(), int(pc.qlen.max()), c.d_lens,
                             int(c.lens.max()), c.sm_scale, True, (-1, -1), c.d_bt, 0.0, None,
                             c.d_ks.reshape(1, -1).expand(R, c.Hkv), c.d_vs.reshape(1, -1).expand(R, c.Hkv),
                             kv_quant_mode=KVQuantMode.FP8_PER_TENSOR)
    except Exception as e:  # noqa: BLE001
        pytest.skip(f"vLLM Triton unified_attention unavailable here: {type(e).__name__}: {e}")
    torch.cuda.synchronize()
    close(to_np_f32(out), to_np_f32(ref), "bf16", "vs vLLM Triton unified_attention")


def test_prefill_rejects_16_bit_caches_and_bad_heads(ops):
    c = Case([64], 8, 1, kv="bf16").cuda()
    q = torch.zeros(4, 8, 128, dtype=torch.bfloat16, device="cuda")
    qsl = torch.tensor([0, 4], dtype=torch.int32, device="cuda")
    tp = torch.zeros(2, dtype=torch.int32, device="cuda")
    with pytest.raises(ops.B200AttnError, match="fp8_e4m3"):
        torch.ops.b200attn.paged_prefill(q, c.kc, c.vc, c.d_bt, c.d_lens, qsl, tp, torch.empty_like(q), c.d_ks, c.d_vs, 0.1, 0)
    with pytest.raises(ops.B200AttnError, match="head counts"):
        ops.prefill_tile_tokens(9, 2)
