"""Golden vectors for the prefill / mixed-batch oracle (oracle.paged_prefill_ref) from the reference's own arithmetic:
vLLM 0.22.0's Triton `unified_attention` ($VLLM/v1/attention/ops/triton_unified_attention.py) with q_len > 1, executed
unmodified on CPU under TRITON_INTERPRET=1 on a small seeded mixed batch (a decode row, chunks with and without cached
context, a chunk longer than the kernel's query tile) over an FP8-e4m3 and an fp16 paged cache.

Same limits as make_golden.py: fp16 activations (the interpreter mishandles bf16), fp8 inputs inside +-448.
Run from the repo root (CPU only, ~1-2 min):   python tests/golden/make_golden_prefill.py
"""
import math
import os
import sys

os.environ["TRITON_INTERPRET"] = "1"
import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
from oracle.oracle import e4m3_encode_satfinite, make_block_table  # noqa: E402

from vllm.v1.attention.ops.triton_unified_attention import unified_attention  # noqa: E402
from vllm.v1.kv_cache_interface import KVQuantMode  # noqa: E402
import vllm  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    torch.manual_seed(20261105)
    Hq, Hkv, D, bs = 8, 2, 128, 16
    ctx = np.array([40, 0, 75, 0], dtype=np.int32)
    qlen = np.array([1, 21, 9, 50], dtype=np.int32)
    lens = ctx + qlen
    R = len(lens)
    bt, nb = make_block_table(lens, bs, seed=123)
    qsl = np.concatenate([[0], np.cumsum(qlen)]).astype(np.int32)
    T = int(qlen.sum())
    q = torch.randn(T, Hq, D, dtype=torch.float16)
    ks = torch.tensor([0.011], dtype=torch.float32)
    vs = torch.tensor([0.013], dtype=torch.float32)
    k16 = torch.randn(nb, bs, Hkv, D, dtype=torch.float16)
    v16 = torch.randn(nb, bs, Hkv, D, dtype=torch.float16)
    k8 = torch.from_numpy(e4m3_encode_satfinite((k16.float() / ks).numpy())).view(torch.float8_e4m3fn)
    v8 = torch.from_numpy(e4m3_encode_satfinite((v16.float() / vs).numpy())).view(torch.float8_e4m3fn)
    sm_scale = 1.0 / math.sqrt(D)

    def run(kc, vc, fp8):
        o = torch.zeros(T, Hq, D, dtype=torch.float16)
        unified_attention(q, kc, vc, o, torch.from_numpy(qsl), int(qlen.max()), torch.from_numpy(lens), int(lens.max()), sm_scale,
                          True, (-1, -1), torch.from_numpy(bt), 0.0, None, ks.expand(R, Hkv) if fp8 else None,
                          vs.expand(R, Hkv) if fp8 else None,
                          kv_quant_mode=KVQuantMode.FP8_PER_TENSOR if fp8 else KVQuantMode.NONE)
        return o.numpy()

    out = {"vllm_version": np.array(vllm.__version__), "ctx": ctx, "qlen": qlen, "lens": lens, "query_start_loc": qsl,
           "block_table": bt, "block_size": np.int32(bs), "q": q.numpy(), "sm_scale": np.float32(sm_scale),
           "k_scale": ks.numpy(), "v_scale": vs.numpy(),
           "k_cache_fp8": k8.view(torch.uint8).numpy(), "v_cache_fp8": v8.view(torch.uint8).numpy(),
           "k_cache_f16": k16.numpy(), "v_cache_f16": v16.numpy()}
    out["out_fp8"] = run(k8, v8, True)
    out["out_f16"] = run(k16, v16, False)
    path = os.path.join(HERE, "vllm_triton_interp_prefill_fp16.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
