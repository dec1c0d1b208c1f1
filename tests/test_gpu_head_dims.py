"""head_dim 64 and 256 (csrc/decode_simt.cu behind b200attn_decode; kv_write / rope_kv_write / kv_gather are head-size
generic): the talker config class has no head_dim field, so class defaults give 1024 / 16 = 64
(vllm_omni/model_executor/models/qwen3_tts/configuration_qwen3_tts.py:376-398; BASELINE configs[0] "D 64 variant"), and
FlashInfer serves 64 / 128 / 256 ($VLLM/v1/attention/backends/flashinfer.py:403-405).  Same oracle, same tolerance rule."""
import numpy as np
import pytest
import torch

from tests.helpers import Case
from tests.test_gpu_parity import run_case

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    assert torch.cuda.is_available()
    import ht_vllm_omni_b200.ops as ops

    ops.load_library()
    return ops


CASES = [
    ("d64_fp8_ragged", [1, 37, 100, 0, 64, 513, 2048], 16, 2, 64, "fp8", None, {}),
    ("d64_bf16_c1_talker_defaults", [2048], 16, 2, 64, "bf16", None, {}),
    ("d64_fp8_per_head_few_ctas", [300, 700, 1500], 16, 2, 64, "fp8", 5, {"per_head": True}),
    ("d64_fp16", [300, 130], 8, 2, 64, "fp16", None, {}),
    ("d64_G16_two_head_passes", [300, 700], 16, 1, 64, "fp8", None, {}),
    ("d64_block64_NHD", [300, 700], 16, 2, 64, "fp8", None, {"bs": 64, "layout": "NHD"}),
    ("d256_fp8_ragged", [1, 37, 100, 64, 513, 1100], 16, 2, 256, "fp8", None, {}),
    ("d256_bf16", [300, 700, 33], 8, 2, 256, "bf16", None, {}),
    ("d256_fp8_long_one_sequence", [9000], 8, 1, 256, "fp8", None, {}),
]


@pytest.mark.parametrize("name,lens,Hq,Hkv,D,kv,ctas,extra", CASES, ids=[c[0] for c in CASES])
def test_write_and_decode_match_oracle(ops, name, lens, Hq, Hkv, D, kv, ctas, extra):
    act = kv if kv != "fp8" else "bf16"
    c = Case(lens, Hq, Hkv, D=D, kv=kv, act=act, **extra).cuda()
    run_case(ops, c, ctas=ctas, lse=True)


def test_lse_and_decode_kv_write_entry(ops):
    """The LSE output and the forward-writes-KV entry point (two launches from one call at these head sizes)."""
    c = Case([5, 300, 1100], 16, 2, D=64, kv="fp8").cuda()
    got, lse, (kc_o, vc_o) = run_case(ops, c, lse=True)
    from oracle import oracle as O

    for s in range(c.B):
        L = int(c.lens[s])
        K = np.stack([O.e4m3_decode(kc_o[c.bt_np[s, p // c.bs], p % c.bs, 0]) for p in range(L)]).astype(np.float64)
        logits = K @ c.q[s, 3].double().numpy() * c.sm_scale * float(c.k_scale[0])
        ref = np.log(np.exp(logits - logits.max()).sum()) + logits.max()
        assert abs(lse[s, 3] - ref) < 2e-3
    # decode_kv_write: history written first, the rows' last tokens by the call itself
    last = np.cumsum(c.lens) - 1
    hist = c.slots_np.copy(); hist[last] = -1
    c.kc.zero_(); c.vc.zero_()
    torch.ops.b200attn.kv_write(c.d_key, c.d_value, c.kc, c.vc, torch.from_numpy(hist).cuda(), c.d_ks, c.d_vs)
    ws = ops.DecodeWorkspace("cuda", c.B, torch.uint8, head_dim=64)
    ws.make_plan(c.d_lens, c.Hq, c.Hkv, c.d_bt, c.bs)
    idx = torch.from_numpy(last).cuda()
    out = torch.zeros(c.B, c.Hq, 64, dtype=torch.bfloat16, device="cuda")
    ws.decode_kv_write(c.d_q, c.d_key[idx].contiguous(), c.d_value[idx].contiguous(), c.d_slots[idx].contiguous(), c.kc, c.vc, c.d_bt,
                       c.d_lens, out, c.d_ks, c.d_vs, c.sm_scale)
    torch.cuda.synchronize()
    assert np.array_equal(c.cache_to_np(c.kc), kc_o) and np.array_equal(c.cache_to_np(c.vc), vc_o)
    assert np.array_equal(out.float().cpu().numpy(), got)


def test_backend_accepts_head_size_64_and_256():
    from ht_vllm_omni_b200 import backend as BK

    be = BK.B200PagedAttentionBackend
    assert be.supports_head_size(64) and be.supports_head_size(256)
    for D in (64, 256):
        BK.B200PagedAttentionImpl(16, D, D ** -0.5, 2, None, None, "fp8")
    with pytest.raises(NotImplementedError):
        BK.B200PagedAttentionImpl(16, 64, 0.1, 2, None, None, BK.PER_TOKEN_HEAD)
