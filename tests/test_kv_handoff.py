"""KV hand-off gather (csrc/kv_gather.cu, ht_vllm_omni_b200/kv_handoff.py).

CPU: oracle.kv_extract_ref is pinned against outputs of the REFERENCE's own `_extract_kv_cache` / `normalize_layer_kv`
(tests/golden/make_golden_kv_extract.py executes their source from /root/reference; fixture tests/golden/ref_kv_extract.npz).
GPU: kv_handoff.gather_request_kv (one launch + one pinned D2H for all layers) reproduces the same fixtures and the oracle
byte for byte -- stacked [nb, 2, ...] / [2, nb, ...] / tuple caches, HND-strided caches, invalid ids, truncation.
"""
import os

import numpy as np
import pytest
import torch

from oracle import oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
CASES = ["fp8_nb2", "bf16_2nb", "fp8_invalid_ids", "bf16_tuple_short"]


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(HERE, "golden", "ref_kv_extract.npz"))


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_extract(golden, name):
    g = golden
    caches = g[name + "_caches"]            # [L, nb, 2, bs, H, D] raw
    for l in range(caches.shape[0]):
        got = O.kv_extract_ref(g[name + "_ids"].tolist(), int(g[name + "_seq_len"]), caches[l][:, 0], caches[l][:, 1])
        assert np.array_equal(got[0], g[name + "_k"][l]) and np.array_equal(got[1], g[name + "_v"][l])
    assert bool(g["all_invalid_is_none"])
    assert O.kv_extract_ref([99, -5], 10, caches[0][:, 0], caches[0][:, 1]) is None


def test_host_logic_rejects_cpu_caches():
    from ht_vllm_omni_b200 import kv_handoff
    from ht_vllm_omni_b200._lib import B200AttnError
    assert kv_handoff.normalize_layer_kv(torch.zeros(3, 3)) is None
    assert kv_handoff.gather_request_kv([0], 4, [torch.zeros(3, 3)]) is None
    with pytest.raises(B200AttnError):
        kv_handoff.gather_request_kv([0], 4, [torch.zeros(4, 2, 16, 2, 64)])   # CPU cache: no fallback


def _raw(t):
    return t.contiguous().view(torch.uint8).numpy() if t.dtype == torch.uint8 else t.contiguous().view(torch.int16).numpy().view(np.uint16)


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_gather_matches_reference_fixture(golden, name):
    from ht_vllm_omni_b200 import kv_handoff
    g = golden
    caches = g[name + "_caches"]
    layout = str(g[name + "_layout"])
    fp8 = caches.dtype == np.uint8
    layer_kv = []
    for l in range(caches.shape[0]):
        t = torch.from_numpy(caches[l].view(np.uint8) if fp8 else caches[l].view(np.int16)).cuda()
        t = t if fp8 else t.view(torch.bfloat16)
        layer_kv.append(t if layout == "nb2" else (t.transpose(0, 1).contiguous() if layout == "2nb" else (t[:, 0], t[:, 1])))
    r = kv_handoff.extract_kv_cache("req", g[name + "_ids"].tolist(), int(g[name + "_seq_len"]), layer_kv, 16, "auto")
    assert r["metadata"]["seq_len"] == int(g[name + "_seq_len"]) and r["metadata"]["num_layers"] == caches.shape[0]
    for l in range(caches.shape[0]):
        assert np.array_equal(_raw(r["layer_blocks"]["key_cache"][l]), g[name + "_k"][l])
        assert np.array_equal(_raw(r["layer_blocks"]["value_cache"][l]), g[name + "_v"][l])
    assert kv_handoff.extract_kv_cache("req", [99, -5], 10, layer_kv, 16, "auto") is None


@pytest.mark.gpu
@pytest.mark.parametrize("kv", ["fp8", "bf16"])
def test_gather_from_hnd_backend_cache_matches_oracle(kv):
    """Our backend's cache: logical [nb, 2, bs, H, D] over HND physical memory, 20 layers, ragged tail, ids out of range."""
    from ht_vllm_omni_b200 import kv_handoff
    nb, bs, H, D, L = 300, 16, 2, 128, 20
    gen = torch.Generator().manual_seed(9)
    dt = torch.uint8 if kv == "fp8" else torch.bfloat16
    caches = []
    for _ in range(L):
        phys = (torch.randint(0, 256, (nb, 2, H, bs, D), generator=gen, dtype=torch.uint8) if kv == "fp8"
                else torch.randn(nb, 2, H, bs, D, generator=gen).to(dt)).cuda()
        caches.append(phys.permute(0, 1, 3, 2, 4))
    ids = torch.randperm(nb, generator=gen)[:257].tolist()
    ids[5], ids[100] = -3, nb + 7
    seq_len = 255 * bs - 9
    got = kv_handoff.gather_request_kv(ids, seq_len, caches)
    for l in (0, 7, 19):
        c = caches[l].cpu()
        c = c.numpy() if kv == "fp8" else c.view(torch.int16).numpy()
        ref = O.kv_extract_ref(ids, seq_len, c[:, 0], c[:, 1])
        k = got[0][l].numpy() if kv == "fp8" else got[0][l].view(torch.int16).numpy()
        v = got[1][l].numpy() if kv == "fp8" else got[1][l].view(torch.int16).numpy()
        assert k.shape == (seq_len, H, D)
        assert np.array_equal(k, ref[0]) and np.array_equal(v, ref[1])


def test_kv_scale_metadata_is_plain_lists_per_layer():
    """`metadata["kv_scales"]`: what the receiving stage needs to read the raw e4m3 bytes (value = byte * scale[head])."""
    from ht_vllm_omni_b200 import kv_handoff as H

    m = H.kv_scale_metadata([(torch.tensor([0.5]), torch.tensor([0.25])), None, (torch.tensor([0.1, 0.2]), torch.tensor([0.3, 0.4])), (1.0, 2.0)])
    assert m["k_scale"] == [[0.5], None, pytest.approx([0.1, 0.2]), [1.0]]
    assert m["v_scale"] == [[0.25], None, pytest.approx([0.3, 0.4]), [2.0]]
    assert H.kv_scale_metadata([None]) == {"k_scale": [None], "v_scale": [None]}


@pytest.mark.gpu
def test_extract_carries_fp8_scales_and_dequantises_to_the_written_values():
    """Hand-off of an FP8 cache: raw bytes + per-head scales in the metadata reproduce, on the receiving side, exactly the
    values the oracle's write produced (e4m3_decode(byte) * scale[head])."""
    from ht_vllm_omni_b200 import kv_handoff as H

    rng = np.random.default_rng(5)
    nb, bs, Hkv, D, L = 12, 16, 2, 128, 3
    phys = [torch.from_numpy(rng.integers(0, 0x78, size=(nb, 2, Hkv, bs, D), dtype=np.uint8)).cuda() for _ in range(L)]
    caches = [p.permute(0, 1, 3, 2, 4) for p in phys]
    scales = [(torch.tensor([0.01 * (l + 1), 0.02 * (l + 1)], device="cuda"), torch.tensor([0.03, 0.04], device="cuda")) for l in range(L)]
    got = H.extract_kv_cache("r", [3, 9, 1], 40, caches, bs, "fp8", {"note": "x"}, kv_scales=scales)
    md = got["metadata"]
    assert md["note"] == "x" and md["seq_len"] == 40 and md["dtype"] == "fp8" and md["num_layers"] == L
    for l in range(L):
        ks = np.asarray(md["kv_scales"]["k_scale"][l], np.float32); vs = np.asarray(md["kv_scales"]["v_scale"][l], np.float32)
        assert np.allclose(ks, [0.01 * (l + 1), 0.02 * (l + 1)]) and np.allclose(vs, [0.03, 0.04])
        c = caches[l].cpu().numpy()
        k_ref, _ = O.kv_extract_ref([3, 9, 1], 40, c[:, 0], c[:, 1])
        k_bytes = got["layer_blocks"]["key_cache"][l].numpy()
        assert np.array_equal(k_bytes, k_ref)
        deq = O.e4m3_decode(k_bytes) * ks[None, :, None]
        assert np.array_equal(deq, O.e4m3_decode(k_ref) * ks[None, :, None])
    # without kv_scales the metadata is exactly the reference's
    plain = H.extract_kv_cache("r", [3, 9, 1], 40, caches, bs, "fp8")
    assert set(plain["metadata"]) == {"block_size", "num_layers", "dtype", "seq_len"}
