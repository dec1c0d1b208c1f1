"""b200attn_decode_kv_write (a backend with forward_includes_kv_cache_update = True): the decode launch that also writes
the step's new tokens must leave the SAME cache bytes as kv_write (bit-exact against the oracle) and produce the SAME
attention output as kv_write followed by decode -- whether the store happens inside the tcgen05 kernel (FP8 cache, <= 8 query
heads per kv head) or in a KV-write launch issued by the same call (every other shape)."""
import numpy as np
import pytest
import torch

from tests.helpers import Case

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import ht_vllm_omni_b200.ops as ops

    ops.load_library()
    return ops


CASES = [
    # name, lens, Hq, Hkv, kv, ctas, extra
    ("L1_only_the_new_token", [1, 1, 1], 16, 2, "fp8", None, {}),
    ("new_token_opens_a_tile", [129, 257, 1025, 4097], 16, 2, "fp8", None, {}),
    ("new_token_closes_a_tile", [128, 256, 1024, 64, 16], 16, 2, "fp8", None, {}),
    ("ragged_with_empty_rows", [1, 37, 100, 0, 64, 513, 0, 2048], 16, 2, "fp8", None, {}),
    ("cut_rows_7_ctas", [300, 700, 1500, 90], 16, 2, "fp8", 7, {"per_head": True}),
    ("cut_rows_cluster_pairs", [1024] * 8, 16, 2, "fp8", 32, {}),
    ("thinker_heads", [777, 1500, 64, 2000], 32, 4, "fp8", None, {"per_head": True}),
    ("fp16_activations", [300, 700], 16, 2, "fp8", 5, {"act": "fp16"}),
    ("block_size_64", [300, 700, 1000], 16, 2, "fp8", None, {"bs": 64}),
    ("G16_two_launches", [300, 700], 16, 1, "fp8", 3, {}),
    ("bf16_cache_two_launches", [300, 700, 33], 16, 2, "bf16", None, {}),
]


def _run(ops, c, ctas, fused_option=1):
    lib = ops.load_library()
    last = np.cumsum(c.lens) - 1                       # index of each row's last token in c.key
    has = c.lens > 0
    slots_hist = c.slots_np.copy()
    slots_hist[last[has]] = -1                         # history: everything but the rows' last tokens
    new_slots = np.full(c.B, -1, np.int64)
    new_slots[has] = c.slots_np[last[has]]
    idx = torch.from_numpy(np.where(has, last, 0)).to("cuda")
    new_k, new_v = c.d_key[idx].contiguous(), c.d_value[idx].contiguous()
    d_new_slots = torch.from_numpy(new_slots).to("cuda")
    kvdt = torch.uint8 if c.fp8 else c.d_q.dtype

    # reference flow: KV write of every token, then decode
    torch.ops.b200attn.kv_write(c.d_key, c.d_value, c.kc, c.vc, c.d_slots, c.d_ks, c.d_vs)
    ws = ops.DecodeWorkspace("cuda", c.B, kvdt, num_ctas=ctas, group_size=c.Hq // c.Hkv)
    ws.make_plan(c.d_lens, c.Hq, c.Hkv, c.d_bt, c.bs)
    out_a = torch.full((c.B, c.Hq, c.D), float("nan"), dtype=c.d_q.dtype, device="cuda")
    ws.decode(c.d_q, c.kc, c.vc, c.d_bt, c.d_lens, out_a, c.d_ks, c.d_vs, c.sm_scale, fresh_tokens=1)
    torch.cuda.synchronize()
    kc_a, vc_a = c.cache_to_np(c.kc), c.cache_to_np(c.vc)

    # fused flow: history first (the earlier steps), then ONE call for the step's write + attention
    c.kc.zero_(); c.vc.zero_()
    torch.ops.b200attn.kv_write(c.d_key, c.d_value, c.kc, c.vc, torch.from_numpy(slots_hist).to("cuda"), c.d_ks, c.d_vs)
    prev = lib.b200attn_set_option(b"fused_write", fused_option)
    try:
        out_b = torch.full((c.B, c.Hq, c.D), float("nan"), dtype=c.d_q.dtype, device="cuda")
        ws.make_plan(c.d_lens, c.Hq, c.Hkv, c.d_bt, c.bs)
        ws.decode_kv_write(c.d_q, new_k, new_v, d_new_slots, c.kc, c.vc, c.d_bt, c.d_lens, out_b, c.d_ks, c.d_vs, c.sm_scale)
        torch.cuda.synchronize()
    finally:
        lib.b200attn_set_option(b"fused_write", prev)
    kc_o, vc_o = c.oracle_cache()
    assert np.array_equal(kc_a, kc_o) and np.array_equal(vc_a, vc_o)
    assert np.array_equal(c.cache_to_np(c.kc), kc_o), "K cache bytes after the fused write differ from the oracle"
    assert np.array_equal(c.cache_to_np(c.vc), vc_o), "V cache bytes after the fused write differ from the oracle"
    a, b = out_a.float().cpu().numpy()[has], out_b.float().cpu().numpy()[has]
    assert not np.isnan(b).any()
    assert np.array_equal(a, b), f"fused-write output differs from write-then-decode: max {np.abs(a - b).max():.3e}"


@pytest.mark.parametrize("name,lens,Hq,Hkv,kv,ctas,extra", CASES, ids=[c[0] for c in CASES])
def test_decode_kv_write_equals_write_then_decode(ops, name, lens, Hq, Hkv, kv, ctas, extra):
    extra = dict(extra)
    act = extra.pop("act", "bf16")
    c = Case(lens, Hq, Hkv, kv=kv, act=act if kv == "fp8" else kv, seed=len(lens), **extra).cuda()
    _run(ops, c, ctas)


def test_decode_kv_write_option_off_uses_two_launches(ops):
    c = Case([300, 700, 1500, 90], 16, 2, kv="fp8", seed=3).cuda()
    _run(ops, c, 7, fused_option=0)


def test_decode_kv_write_repeated_steps_streaming(ops):
    """Eight consecutive decode steps of a growing batch, each step's token written by the decode launch itself: after the
    last step the cache equals the oracle's and every step's output equals write-then-decode on a second cache."""
    steps, B, Hq, Hkv = 8, 6, 16, 2
    base = np.array([100, 127, 128, 250, 511, 1000], np.int32)
    c = Case((base + steps).tolist(), Hq, Hkv, kv="fp8", seed=11, per_head=True).cuda()
    starts = np.concatenate([[0], np.cumsum(c.lens)[:-1]])
    hist = np.full_like(c.slots_np, -1)
    for r in range(B):
        hist[starts[r]:starts[r] + base[r]] = c.slots_np[starts[r]:starts[r] + base[r]]
    d_hist = torch.from_numpy(hist).to("cuda")
    _, kc2, vc2 = __import__("tests.helpers", fromlist=["alloc_cache"]).alloc_cache(c.nb, c.bs, Hkv, c.D, "fp8")
    torch.ops.b200attn.kv_write(c.d_key, c.d_value, c.kc, c.vc, d_hist, c.d_ks, c.d_vs)
    torch.ops.b200attn.kv_write(c.d_key, c.d_value, kc2, vc2, d_hist, c.d_ks, c.d_vs)
    ws = ops.DecodeWorkspace("cuda", B, torch.uint8)
    for t in range(steps):
        lens = torch.from_numpy(base + t + 1).to("cuda")
        idx = torch.from_numpy(starts + base + t).to("cuda")
        k, v, sl = c.d_key[idx].contiguous(), c.d_value[idx].contiguous(), c.d_slots[idx].contiguous()
        out_a = torch.zeros(B, Hq, c.D, dtype=torch.bfloat16, device="cuda")
        out_b = torch.zeros_like(out_a)
        ws.make_plan(lens, Hq, Hkv, c.d_bt, c.bs)
        ws.decode_kv_write(c.d_q, k, v, sl, c.kc, c.vc, c.d_bt, lens, out_b, c.d_ks, c.d_vs, c.sm_scale)
        torch.ops.b200attn.kv_write(k, v, kc2, vc2, sl, c.d_ks, c.d_vs)
        ws.make_plan(lens, Hq, Hkv, c.d_bt, c.bs)
        ws.decode(c.d_q, kc2, vc2, c.d_bt, lens, out_a, c.d_ks, c.d_vs, c.sm_scale, fresh_tokens=1)
        torch.cuda.synchronize()
        assert torch.equal(out_a, out_b), f"step {t}"
    kc_o, vc_o = c.oracle_cache()
    assert np.array_equal(c.cache_to_np(c.kc), kc_o) and np.array_equal(c.cache_to_np(c.vc), vc_o)
