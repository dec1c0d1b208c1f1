"""CPU: pin the oracle (oracle/oracle.py) against golden vectors produced by the reference's own arithmetic
(vLLM 0.22.0 Triton kernels run under TRITON_INTERPRET=1, tests/golden/make_golden.py)."""
import numpy as np
import pytest
import torch

from oracle import oracle as O


def test_e4m3_codec_matches_torch_in_range():
    # every finite e4m3 value, its neighbours' midpoints (ties) and random in-range floats
    vals = O.E4M3_LUT[np.isfinite(O.E4M3_LUT)]
    mids = (np.sort(vals)[:-1] + np.sort(vals)[1:]) / 2
    rng = np.random.default_rng(0)
    x = np.concatenate([vals, mids, rng.uniform(-448, 448, 20000), rng.normal(0, 1e-2, 20000)]).astype(np.float32)
    ours = O.e4m3_encode_satfinite(x)
    ref = torch.from_numpy(x).to(torch.float8_e4m3fn).view(torch.uint8).numpy()
    # torch encodes -0.0 as 0x80 too; compare bytes exactly
    assert np.array_equal(ours, ref)
    assert np.array_equal(O.e4m3_decode(ours), torch.from_numpy(ref).view(torch.float8_e4m3fn).float().numpy())


def test_e4m3_saturates_and_keeps_nan():
    x = np.array([449.0, 464.0, 1e9, np.inf, -1e9, -np.inf, np.nan, 448.0, -448.0], dtype=np.float32)
    got = O.e4m3_encode_satfinite(x)
    assert list(got[:6]) == [0x7E, 0x7E, 0x7E, 0x7E, 0xFE, 0xFE]   # __NV_SATFINITE: clamp to +-448
    assert got[6] & 0x7F == 0x7F and got[7] == 0x7E and got[8] == 0xFE


def test_bf16_round_matches_torch():
    rng = np.random.default_rng(1)
    x = np.concatenate([rng.normal(0, 1, 50000), rng.normal(0, 1e-20, 100), [0.0, -0.0, 1.00390625, 3.3895e38]]).astype(np.float32)
    ref = torch.from_numpy(x).to(torch.bfloat16).float().numpy()
    assert np.array_equal(O.bf16_round(x), ref)


def _assert_fp8_equal_modulo_interpreter_cast_defects(ours, ref, quot):
    """Bytes must be identical, except for two documented defects of the Triton *interpreter's* software fp32->fp8
    cast (not of the reference algorithm; absent on hardware and in torch's CPU cast, which both agree with the oracle
    -- see test_e4m3_codec_matches_torch_in_range, and tests/test_gpu_vs_vllm.py for vLLM's CUDA op on the GPU box):
      (a) dropped carry: when rounding carries the mantissa into the next binade (127.15 -> 128) the interpreter
          stores 2^e (64) instead of 2^(e+1);
      (b) exact ties are rounded away from zero (100.0 -> 104) instead of to even (96).
    `quot` holds the float32 quotient x/scale at each cache position so (b) can be verified as a true tie."""
    bad = ours != ref
    o, r, qv = ours[bad].astype(np.int32), ref[bad].astype(np.int32), quot[bad].astype(np.float64)
    lo, hi = np.abs(O.e4m3_decode(ours[bad])).astype(np.float64), np.abs(O.e4m3_decode(ref[bad])).astype(np.float64)
    carry = ((o & 7) == 0) & (hi == lo / 2)      # ours = 2^(e+1), interpreter = 2^e (also at the subnormal edge)
    tie = (r - o == 1) & (np.abs(qv) == (lo + hi) / 2) & ((o & 1) == 0)
    assert np.all(carry | tie), "mismatch outside the two known interpreter cast defects"
    assert bad.mean() < 0.03


def test_kv_write_fp8_bit_exact_vs_reference_kernel(golden):
    g = golden
    key, value = g["key"].astype(np.float32), g["value"].astype(np.float32)
    kc = np.zeros_like(g["k_cache_fp8"]); vc = np.zeros_like(g["v_cache_fp8"])
    O.kv_write_ref(key, value, kc, vc, g["slots"], g["k_scale"], g["v_scale"], fp8=True)
    kq = np.zeros(kc.shape, np.float32); vq = np.zeros(vc.shape, np.float32)
    O.kv_write_ref(key / g["k_scale"][0], value / g["v_scale"][0], kq, vq, g["slots"], fp8=False)
    _assert_fp8_equal_modulo_interpreter_cast_defects(kc, g["k_cache_fp8"], kq)
    _assert_fp8_equal_modulo_interpreter_cast_defects(vc, g["v_cache_fp8"], vq)
    # untouched slots (never written, incl. the padding token's) stay zero in both
    assert np.array_equal(kc == 0, g["k_cache_fp8"] == 0)
    assert (g["slots"] < 0).any()  # the fixture holds a padding token


def test_kv_write_16bit_bit_exact_vs_reference_kernel(golden):
    g = golden
    kc = np.zeros(g["k_cache_f16"].shape, np.float32); vc = np.zeros(g["v_cache_f16"].shape, np.float32)
    O.kv_write_ref(g["key"].astype(np.float32), g["value"].astype(np.float32), kc, vc, g["slots"], fp8=False)
    assert np.array_equal(kc.astype(np.float16), g["k_cache_f16"])
    assert np.array_equal(vc.astype(np.float16), g["v_cache_f16"])


@pytest.mark.parametrize("kind", ["fp8", "f16"])
def test_decode_matches_reference_2d_kernel(golden, kind):
    g = golden
    fp8 = kind == "fp8"
    kc = g["k_cache_fp8"] if fp8 else g["k_cache_f16"].astype(np.float32)
    vc = g["v_cache_fp8"] if fp8 else g["v_cache_f16"].astype(np.float32)
    out = O.paged_decode_ref(g["q"].astype(np.float32), kc, vc, g["block_table"], g["lens"], float(g["sm_scale"]),
                             g["k_scale"], g["v_scale"], fp8=fp8, act="fp16", tile=16)
    ref = g[f"out_{kind}_2d"].astype(np.float32)
    # same algorithm, same rounding points; the residue is dot-product summation order -> <= 1 fp16 ulp at |x|<4
    assert np.abs(out - ref).max() <= 1e-3
    assert np.mean(out == ref) > 0.8


@pytest.mark.parametrize("kind", ["fp8", "f16"])
def test_split_decode_matches_reference_3d_kernel(golden, kind):
    g = golden
    fp8 = kind == "fp8"
    kc = g["k_cache_fp8"] if fp8 else g["k_cache_f16"].astype(np.float32)
    vc = g["v_cache_fp8"] if fp8 else g["v_cache_f16"].astype(np.float32)
    out = O.paged_decode_split_ref(g["q"].astype(np.float32), kc, vc, g["block_table"], g["lens"], float(g["sm_scale"]),
                                   g["k_scale"], g["v_scale"], fp8=fp8, act="fp16", tile=16, num_segments=4)
    ref = g[f"out_{kind}_3d"].astype(np.float32)
    assert np.abs(out - ref).max() <= 1e-3
    assert np.mean(out == ref) > 0.8


def test_reference_style_oracle_close_to_exact_definition(golden):
    g = golden
    ex = O.paged_decode_exact(g["q"].astype(np.float32), g["k_cache_fp8"], g["v_cache_fp8"], g["block_table"], g["lens"],
                              float(g["sm_scale"]), g["k_scale"], g["v_scale"], fp8=True)
    out = O.paged_decode_ref(g["q"].astype(np.float32), g["k_cache_fp8"], g["v_cache_fp8"], g["block_table"], g["lens"],
                             float(g["sm_scale"]), g["k_scale"], g["v_scale"], fp8=True, act="bf16")
    assert np.abs(out - ex).max() < 1e-2      # BASELINE.json tolerance for FP8 KV


def test_merge_lse_ref_is_consistent_with_whole_attention():
    rng = np.random.default_rng(3)
    n, Hq, D, P = 3, 4, 16, 3
    s = rng.normal(size=(P, n, Hq, 20)).astype(np.float32)   # logits per part
    v = rng.normal(size=(P, n, Hq, 20, D)).astype(np.float32)
    parts_o = np.stack([(np.exp(s[p] - s[p].max(-1, keepdims=True))[..., None] * v[p]).sum(-2) /
                        np.exp(s[p] - s[p].max(-1, keepdims=True)).sum(-1)[..., None] for p in range(P)])
    parts_lse = np.stack([s[p].max(-1) + np.log(np.exp(s[p] - s[p].max(-1, keepdims=True)).sum(-1)) for p in range(P)])
    o, lse = O.merge_lse_ref(parts_o, parts_lse)
    S = np.concatenate(list(s), axis=-1); Vv = np.concatenate(list(v), axis=-2)
    w = np.exp(S - S.max(-1, keepdims=True)); full = (w[..., None] * Vv).sum(-2) / w.sum(-1)[..., None]
    assert np.allclose(o, full, atol=1e-5)
    assert np.allclose(lse, S.max(-1) + np.log(w.sum(-1)), atol=1e-5)


@pytest.fixture(scope="module")
def golden_prefill():
    import os

    path = os.path.join(os.path.dirname(__file__), "golden", "vllm_triton_interp_prefill_fp16.npz")
    return dict(np.load(path))


@pytest.mark.parametrize("kind", ["fp8", "f16"])
def test_prefill_oracle_matches_reference_kernel_on_a_mixed_batch(golden_prefill, kind):
    """oracle.paged_prefill_ref against vLLM's unified_attention with q_len > 1 (tests/golden/make_golden_prefill.py): a decode
    row, chunks with and without cached context, a chunk longer than the kernel's query tile."""
    g = golden_prefill
    fp8 = kind == "fp8"
    kc = g["k_cache_fp8"] if fp8 else g["k_cache_f16"].astype(np.float32)
    vc = g["v_cache_fp8"] if fp8 else g["v_cache_f16"].astype(np.float32)
    out = O.paged_prefill_ref(g["q"].astype(np.float32), g["query_start_loc"], kc, vc, g["block_table"], g["lens"],
                              float(g["sm_scale"]), g["k_scale"], g["v_scale"], fp8=fp8, act="fp16")
    ref = g[f"out_{kind}"].astype(np.float32)
    assert np.abs(ref).max() > 0.1
    assert np.abs(out - ref).max() <= 1e-3           # residue: fp32 summation order (the kernel tiles the row)
    assert np.mean(out == ref) > 0.8
    exact = O.paged_prefill_ref(g["q"].astype(np.float32), g["query_start_loc"], kc, vc, g["block_table"], g["lens"],
                                float(g["sm_scale"]), g["k_scale"], g["v_scale"], fp8=fp8, act="fp16", exact=True)
    assert np.abs(exact - ref).max() <= 2e-3
    # sampled rows agree with the full evaluation
    rows = [0, 5, 30, 80]
    part = O.paged_prefill_ref(g["q"].astype(np.float32), g["query_start_loc"], kc, vc, g["block_table"], g["lens"],
                               float(g["sm_scale"]), g["k_scale"], g["v_scale"], fp8=fp8, act="fp16", rows=rows)
    assert np.abs(part[rows] - out[rows]).max() <= 1e-3   # (BLAS blocks a 4-row product differently: last-ulp residue)
