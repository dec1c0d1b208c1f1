"""Generate golden vectors for the oracle from the reference's own arithmetic.

The reference (ht-vllm-omni) delegates this path to vLLM; vLLM's Triton kernels are importable in this container
(vllm 0.22.0) and run on CPU under TRITON_INTERPRET=1.  We execute, unmodified:
  * reshape_and_cache_kernel_flash            ($VLLM/v1/attention/ops/triton_reshape_and_cache_flash.py:17-124)
  * unified_attention (2D and 3D+reduce_segments)  ($VLLM/v1/attention/ops/triton_unified_attention.py:763-1058)
on small seeded inputs and store inputs + outputs in tests/golden/*.npz.

Limits of this pin (stated in DESIGN.md): activations are fp16 because the Triton interpreter mis-handles bf16
(it yields 1e8-scale garbage); fp8 inputs are kept inside +-448 because the interpreter's out-of-range fp32->fp8
cast is not the hardware's satfinite cast.  The kernel launcher's `torch.cuda.get_device_capability` call is
bypassed by launching the jitted kernel directly with the launcher's own argument recipe.

Run from the repo root (CPU only, ~1 min):   python tests/golden/make_golden.py
"""
import math
import os
import sys

os.environ["TRITON_INTERPRET"] = "1"
import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
from oracle.oracle import make_block_table, slots_for_positions  # noqa: E402

from vllm.v1.attention.ops.triton_reshape_and_cache_flash import reshape_and_cache_kernel_flash  # noqa: E402
from vllm.v1.attention.ops.triton_unified_attention import unified_attention  # noqa: E402
from vllm.v1.kv_cache_interface import KVQuantMode  # noqa: E402
import vllm  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def ref_kv_write(key, value, kc, vc, slots, ks, vs, fp8):
    T, H, D = key.shape
    n = H * D
    tile = min(2048, 1 << (n - 1).bit_length())
    grid = (T, -(-n // tile))
    reshape_and_cache_kernel_flash[grid](
        key_ptr=key, value_ptr=value, key_cache_ptr=kc, value_cache_ptr=vc, slot_mapping_ptr=slots, k_scale=ks, v_scale=vs,
        key_stride=key.stride(0), value_stride=value.stride(0), block_stride=kc.stride(0), head_stride=kc.stride(2),
        dim_stride_k=0, dim_stride_v=0, page_stride=kc.stride(1), num_heads=H, head_size=D, block_size=kc.shape[1], x=1,
        USE_HEAD_MAJOR_LAYOUT=False, FP8_KV_CACHE=fp8, TILE_SIZE=tile)


def main():
    torch.manual_seed(20261104)
    B, Hq, Hkv, D, bs = 4, 8, 2, 128, 16
    lens = np.array([1, 37, 100, 64], dtype=np.int32)
    bt, nb = make_block_table(lens, bs, seed=99)
    # every token of every sequence is written through the reference write kernel
    positions = [(r, p) for r in range(B) for p in range(int(lens[r]))]
    positions.insert(5, (-1, 0))            # a padding token (slot -1) in the middle
    slots = slots_for_positions(bt, positions, bs)
    T = len(positions)
    key = torch.randn(T, Hkv, D, dtype=torch.float16)
    value = torch.randn(T, Hkv, D, dtype=torch.float16)
    key[7, 0, :4] = torch.tensor([0.0, -0.0, 6e-5, -3.0], dtype=torch.float16)   # zero, -0, tiny, plain
    out = {"vllm_version": np.array(vllm.__version__), "lens": lens, "block_table": bt, "slots": slots,
           "key": key.numpy(), "value": value.numpy(), "block_size": np.int32(bs), "Hq": np.int32(Hq)}
    # per-tensor scales chosen so |x/scale| stays below 448 (interpreter cast limit, see module docstring)
    ks = torch.tensor([float(key.abs().max()) / 440.0], dtype=torch.float32)
    vs = torch.tensor([float(value.abs().max()) / 300.0], dtype=torch.float32)
    out["k_scale"], out["v_scale"] = ks.numpy(), vs.numpy()

    kc8 = torch.zeros(nb, bs, Hkv, D, dtype=torch.uint8).view(torch.float8_e4m3fn)
    vc8 = torch.zeros(nb, bs, Hkv, D, dtype=torch.uint8).view(torch.float8_e4m3fn)
    ref_kv_write(key, value, kc8, vc8, torch.from_numpy(slots), ks, vs, True)
    out["k_cache_fp8"], out["v_cache_fp8"] = kc8.view(torch.uint8).numpy(), vc8.view(torch.uint8).numpy()

    kc16 = torch.zeros(nb, bs, Hkv, D, dtype=torch.float16)
    vc16 = torch.zeros(nb, bs, Hkv, D, dtype=torch.float16)
    ref_kv_write(key, value, kc16, vc16, torch.from_numpy(slots), ks, vs, False)
    out["k_cache_f16"], out["v_cache_f16"] = kc16.numpy(), vc16.numpy()

    q = torch.randn(B, Hq, D, dtype=torch.float16)
    out["q"] = q.numpy()
    sm_scale = 1.0 / math.sqrt(D)
    out["sm_scale"] = np.float32(sm_scale)
    cu = torch.arange(B + 1, dtype=torch.int32)
    sl = torch.from_numpy(lens)
    btt = torch.from_numpy(bt)

    def run(kc, vc, fp8, segs):
        o = torch.zeros(B, Hq, D, dtype=torch.float16)
        kw = {}
        if segs:
            kw = dict(seq_threshold_3D=128, num_par_softmax_segments=segs,
                      softmax_segm_output=torch.zeros(B, Hq, segs, D, dtype=torch.float32),
                      softmax_segm_max=torch.zeros(B, Hq, segs, dtype=torch.float32),
                      softmax_segm_expsum=torch.zeros(B, Hq, segs, dtype=torch.float32))
        unified_attention(q, kc, vc, o, cu, 1, sl, int(lens.max()), sm_scale, True, (-1, -1), btt, 0.0, None,
                          ks.expand(B, Hkv) if fp8 else None, vs.expand(B, Hkv) if fp8 else None,
                          kv_quant_mode=KVQuantMode.FP8_PER_TENSOR if fp8 else KVQuantMode.NONE, **kw)
        return o.numpy()

    out["out_fp8_2d"] = run(kc8, vc8, True, 0)
    out["out_fp8_3d"] = run(kc8, vc8, True, 4)
    out["out_f16_2d"] = run(kc16, vc16, False, 0)
    out["out_f16_3d"] = run(kc16, vc16, False, 4)
    path = os.path.join(HERE, "vllm_triton_interp_fp16.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
