"""Golden vectors for the fused RoPE + KV-write row from vLLM's own arithmetic.

Runs, unmodified and on CPU, vLLM 0.22's RotaryEmbedding code:
  * RotaryEmbeddingBase._compute_inv_freq / _compute_cos_sin_cache   ($VLLM/model_executor/layers/rotary_embedding/base.py)
  * RotaryEmbeddingBase.forward_static -> ApplyRotaryEmb.forward_static  (base.py:130-180, common.py:144-183)
on seeded bf16 / fp16 inputs, neox and GPT-J style, full and partial rotary_dim, and stores inputs + outputs as raw
16-bit patterns in tests/golden/vllm_rope_native.npz.  The instance is made with object.__new__ because the class's
__init__ insists on a process-wide vLLM config; the methods executed are vLLM's.

Run from the repo root (CPU only, seconds):   python tests/golden/make_golden_rope.py
"""
import os

import numpy as np
import torch
import vllm
from vllm.model_executor.layers.rotary_embedding.base import RotaryEmbedding

HERE = os.path.dirname(os.path.abspath(__file__))


def bits(t: torch.Tensor) -> np.ndarray:
    return t.contiguous().view(torch.int16).numpy().view(np.uint16).copy()


def vllm_cache(head_size, rotary_dim, max_pos, base, is_neox, dtype):
    r = object.__new__(RotaryEmbedding)
    r.head_size, r.rotary_dim, r.max_position_embeddings, r.base = head_size, rotary_dim, max_pos, base
    r.is_neox_style, r.dtype = is_neox, dtype
    return RotaryEmbedding._compute_cos_sin_cache(r).to(dtype)


def main():
    torch.manual_seed(20261104)
    out = {"vllm_version": np.array(vllm.__version__)}
    cases = [("bf16_neox", torch.bfloat16, True, 128), ("bf16_gptj", torch.bfloat16, False, 128),
             ("fp16_neox", torch.float16, True, 128), ("bf16_neox_rot64", torch.bfloat16, True, 64)]
    T, Hq, Hkv, D, max_pos, base = 9, 8, 2, 128, 1100, 1000000.0
    for name, dt, neox, rot in cases:
        cache = vllm_cache(D, rot, max_pos, base, neox, dt)
        pos = torch.tensor([0, 1, 2, 17, 255, 1024, 1099, 33, 7], dtype=torch.int64)
        q = torch.randn(T, Hq * D).to(dt)
        k = torch.randn(T, Hkv * D).to(dt)
        q[3, :4] = torch.tensor([0.0, -0.0, 6e-5, -3.0]).to(dt)
        q2, k2 = RotaryEmbedding.forward_static(pos, q.clone(), k.clone(), D, rot, cache, neox)
        out.update({f"{name}_cache": bits(cache), f"{name}_pos": pos.numpy(), f"{name}_q": bits(q), f"{name}_k": bits(k),
                    f"{name}_q_out": bits(q2), f"{name}_k_out": bits(k2),
                    f"{name}_meta": np.array([T, Hq, Hkv, D, rot, int(neox), max_pos], dtype=np.int64)})
    out["base"] = np.array(base)
    np.savez_compressed(os.path.join(HERE, "vllm_rope_native.npz"), **out)
    print("wrote", os.path.join(HERE, "vllm_rope_native.npz"))


if __name__ == "__main__":
    main()
