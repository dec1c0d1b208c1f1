"""Golden vectors for the q/k RMSNorm + M-RoPE row (VERDICT r01 item 7 / SURVEY 8(f) rank 1) from vLLM's own arithmetic, on CPU:
  * MRotaryEmbedding.forward_native ($VLLM/model_executor/layers/rotary_embedding/mrope.py:263-322) with [3, T] positions,
    chunked and interleaved sections (the Qwen3-Omni thinker uses mrope_section [24, 20, 20], interleaved);
  * the native rms_norm RMSNorm.forward_native dispatches to ($VLLM/ir/ops/layernorm.py:9-21, reached from
    $VLLM/model_executor/layers/layernorm.py:82-93) per head (Qwen3 q_norm / k_norm, qwen3.py:153-158).
Instances are made with object.__new__ (the constructors want a process-wide vLLM config); the methods executed are vLLM's.

Run from the repo root (CPU only, seconds):   python tests/golden/make_golden_mrope_norm.py
"""
import os

import numpy as np
import torch
import vllm
from vllm.ir.ops import layernorm as ir_layernorm
from vllm.model_executor.layers.rotary_embedding.base import RotaryEmbedding
from vllm.model_executor.layers.rotary_embedding.common import ApplyRotaryEmb
from vllm.model_executor.layers.rotary_embedding.mrope import MRotaryEmbedding

HERE = os.path.dirname(os.path.abspath(__file__))


def bits(t: torch.Tensor) -> np.ndarray:
    return t.contiguous().view(torch.int16).numpy().view(np.uint16).copy()


def make_mrope(head_size, rotary_dim, max_pos, base, is_neox, dtype, section, interleaved):
    r = object.__new__(MRotaryEmbedding)
    torch.nn.Module.__init__(r)
    r.head_size, r.rotary_dim, r.max_position_embeddings, r.base = head_size, rotary_dim, max_pos, base
    r.is_neox_style, r.dtype = is_neox, dtype
    r.mrope_section, r.mrope_interleaved = list(section), interleaved
    r.scaling_factor = None
    r.cos_sin_cache = RotaryEmbedding._compute_cos_sin_cache(r).to(dtype)
    r._match_cos_sin_cache_dtype = lambda q: r.cos_sin_cache
    a = object.__new__(ApplyRotaryEmb)
    torch.nn.Module.__init__(a)
    a.is_neox_style, a.enable_fp32_compute = is_neox, False
    r.apply_rotary_emb = a
    return r


def main():
    torch.manual_seed(20261105)
    out = {"vllm_version": np.array(vllm.__version__)}
    T, Hq, Hkv, D, max_pos, base = 11, 8, 2, 128, 1300, 1000000.0
    for name, dt, section, inter in [("bf16_interleaved_24_20_20", torch.bfloat16, (24, 20, 20), True),
                                      ("bf16_chunked_16_24_24", torch.bfloat16, (16, 24, 24), False),
                                      ("fp16_interleaved_24_20_20", torch.float16, (24, 20, 20), True)]:
        r = make_mrope(D, D, max_pos, base, True, dt, section, inter)
        pos = torch.stack([torch.randint(0, max_pos, (T,)), torch.randint(0, max_pos, (T,)), torch.randint(0, max_pos, (T,))]).to(torch.int64)
        pos[:, 0] = torch.tensor([5, 5, 5])          # a text token: all axes equal
        q = torch.randn(T, Hq * D).to(dt)
        k = torch.randn(T, Hkv * D).to(dt)
        q2, k2 = MRotaryEmbedding.forward_native(r, pos, q.clone(), k.clone())
        out.update({f"mrope_{name}_cache": bits(r.cos_sin_cache), f"mrope_{name}_pos": pos.numpy(), f"mrope_{name}_q": bits(q),
                    f"mrope_{name}_k": bits(k), f"mrope_{name}_q_out": bits(q2), f"mrope_{name}_k_out": bits(k2),
                    f"mrope_{name}_meta": np.array([T, Hq, Hkv, D, D, *section, int(inter), max_pos], dtype=np.int64)})
    for name, dt in [("bf16", torch.bfloat16), ("fp16", torch.float16)]:
        x = (torch.randn(T, Hq, D) * 3.0).to(dt)
        w = (1.0 + 0.2 * torch.randn(D)).to(dt)
        eps = 1e-6
        native = ir_layernorm.rms_norm.impls["native"].impl_fn   # the python definition in $VLLM/ir/ops/layernorm.py:9-21
        y = native(x.clone(), w, eps)
        out.update({f"rms_{name}_x": bits(x), f"rms_{name}_w": bits(w), f"rms_{name}_y": bits(y), f"rms_{name}_eps": np.array(eps)})
    np.savez_compressed(os.path.join(HERE, "vllm_mrope_rmsnorm_native.npz"), **out)
    print("wrote", os.path.join(HERE, "vllm_mrope_rmsnorm_native.npz"))


if __name__ == "__main__":
    main()
