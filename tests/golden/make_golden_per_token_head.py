"""Golden vectors for the per-(token, head) dynamic FP8 KV write from vLLM's own kernel: runs, unmodified
and on CPU under TRITON_INTERPRET=1, `_reshape_cache_per_token_head`
($VLLM/v1/attention/ops/triton_reshape_and_cache_flash.py:138-236) with the launcher's own argument recipe (:239-300) on
seeded fp16 inputs (the interpreter mis-handles bf16, see make_golden.py), NHD cache, and stores inputs, cache bytes and scale
caches in tests/golden/vllm_triton_interp_per_token_head.npz.

Run from the repo root (CPU only, ~1 min):   python tests/golden/make_golden_per_token_head.py
"""
import os
import sys

os.environ["TRITON_INTERPRET"] = "1"
import numpy as np
import torch
import triton

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import vllm  # noqa: E402
from vllm.v1.attention.ops.triton_reshape_and_cache_flash import _reshape_cache_per_token_head  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    torch.manual_seed(20261106)
    T, Hkv, D, bs, nb = 23, 2, 128, 16, 4
    key = (torch.randn(T, Hkv, D) * 1.7).to(torch.float16)
    value = (torch.randn(T, Hkv, D) * 0.6).to(torch.float16)
    key[4, 1] = 0.0                      # an all-zero head: scale clamps to 1e-6
    value[9, 0, 3] = 300.0               # an outlier sets the head's scale
    slots = torch.randperm(nb * bs)[:T].to(torch.int64)
    slots[6] = -1
    kc = torch.zeros(nb, bs, Hkv, D, dtype=torch.float8_e4m3fn)
    vc = torch.zeros_like(kc)
    ksc = torch.ones(nb, bs, Hkv, dtype=torch.float32)
    vsc = torch.ones(nb, bs, Hkv, dtype=torch.float32)
    _reshape_cache_per_token_head[(T, Hkv)](
        key_ptr=key, value_ptr=value, key_cache_ptr=kc, value_cache_ptr=vc, k_scale_cache_ptr=ksc, v_scale_cache_ptr=vsc,
        slot_mapping_ptr=slots, stride_key_tok=key.stride(0), stride_key_head=key.stride(1), stride_val_tok=value.stride(0),
        stride_val_head=value.stride(1), stride_kc_blk=kc.stride(0), stride_kc_slot=kc.stride(1), stride_kc_head=kc.stride(2),
        stride_vc_blk=vc.stride(0), stride_vc_slot=vc.stride(1), stride_vc_head=vc.stride(2), stride_ks_blk=ksc.stride(0),
        stride_ks_slot=ksc.stride(1), stride_ks_head=ksc.stride(2), stride_vs_blk=vsc.stride(0), stride_vs_slot=vsc.stride(1),
        stride_vs_head=vsc.stride(2), block_size=bs, head_size=D, head_size_v=D, HEAD_SIZE_PADDED=triton.next_power_of_2(D),
        QUANT_MAX=448.0, QUANT_MIN=-448.0)
    out = {"vllm_version": np.array(vllm.__version__), "key": key.numpy(), "value": value.numpy(), "slots": slots.numpy(),
           "k_cache": kc.view(torch.uint8).numpy(), "v_cache": vc.view(torch.uint8).numpy(), "k_scale_cache": ksc.numpy(),
           "v_scale_cache": vsc.numpy(), "meta": np.array([T, Hkv, D, bs, nb], dtype=np.int64)}
    np.savez_compressed(os.path.join(HERE, "vllm_triton_interp_per_token_head.npz"), **out)
    print("wrote", os.path.join(HERE, "vllm_triton_interp_per_token_head.npz"))


if __name__ == "__main__":
    main()
