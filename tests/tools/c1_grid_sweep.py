"""C1 (BASELINE configs[0]: talker B 1, 2k context, bf16 KV, 16:2 heads) is latency-bound: which decode grid / merge route is fastest?
CUDA graph of 20 launches over 20 layer caches, CUDA events.   python tests/tools/c1_grid_sweep.py [B L kv]"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import torch
import ht_vllm_omni_b200.ops as ops
from tests.tools.sm100_reference import _graph_time

lib = ops.load_library()
dev = torch.device("cuda")
B, L, kv = 1, 2048, "bf16"
if len(sys.argv) > 3:
    B, L, kv = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
Hq, Hkv, D, bs, NL = 16, 2, 128, 16, 20
nblk = -(-L // bs); nb = B * nblk + 8
g = torch.Generator(device=dev).manual_seed(1)
bt = torch.randperm(nb, generator=torch.Generator().manual_seed(7))[: B * nblk].to(torch.int32).reshape(B, nblk).to(dev)
lens = torch.full((B,), L, dtype=torch.int32, device=dev)
fp8 = kv == "fp8"
caches = []
for _ in range(NL):
    if fp8:
        ph = torch.randint(0, 0x78, (nb, 2, Hkv, bs, D), dtype=torch.uint8, device=dev, generator=g)
    else:
        ph = torch.randn((nb, 2, Hkv, bs, D), device=dev, generator=g).to(torch.bfloat16)
    caches.append(ph.permute(0, 1, 3, 2, 4).unbind(1))
q = torch.randn(B, Hq, D, device=dev, generator=g).to(torch.bfloat16)
out = torch.zeros(NL, B, Hq, D, dtype=torch.bfloat16, device=dev)
ks = torch.full((Hkv,), 0.01, device=dev); vs = torch.full((Hkv,), 0.01, device=dev)
kvt = torch.uint8 if fp8 else torch.bfloat16
ref = None
for ctas in (None, 2, 4, 8, 16, 32, 64, 128):
    for fused in (True, False):
        ws = ops.DecodeWorkspace(dev, B, kvt, num_ctas=ctas, group_size=Hq // Hkv)
        ws.make_plan(lens, Hq, Hkv, bt, bs)

        def body():
            for l in range(NL):
                if ctas is None and fused:
                    ws.decode(q, caches[l][0], caches[l][1], bt, lens, out[l], ks, vs, D ** -0.5, fresh_tokens=1)   # library routing
                else:
                    ops.launch_split_decode(q, caches[l][0], caches[l][1], bt, lens, out[l], ws.plan, ws.workspace, ks, vs, D ** -0.5, ws.grid, None, None, 1, fused)
                    if not fused:
                        ops.launch_merge(caches[l][0], lens, out[l], ws.plan, ws.workspace, vs, ws.grid)
        us, _ = _graph_time(body, NL)
        torch.cuda.synchronize()
        if ref is None:
            ref = out.clone()
        err = float((out.float() - ref.float()).abs().max())
        tag = "library default" if (ctas is None and fused) else ("in-kernel merge" if fused else "merge kernel")
        print(f"grid {ws.grid:4d} {tag:16s}: {us:6.2f} us per layer   max|diff| {err:.1e}", flush=True)
