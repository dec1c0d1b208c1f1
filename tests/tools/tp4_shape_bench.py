"""One rank of the thinker at tp 4 (B 32, 32k, 8 query heads : 1 kv head, fp8): decode launch time for grid / cluster / merge variants.
CUDA graph of 4 launches over 4 distinct caches, CUDA events.   python tests/tools/tp4_shape_bench.py"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import torch
import ht_vllm_omni_b200.ops as ops
from tests.tools.sm100_reference import _graph_time

lib = ops.load_library()
dev = torch.device("cuda")
B, L, Hq, Hkv, D, bs, NL = 32, 32768, 8, 1, 128, 16, 4
if len(sys.argv) > 1:
    B, L, Hq, Hkv = (int(x) for x in sys.argv[1:5])
nblk = L // bs; nb = B * nblk + 8
g = torch.Generator(device=dev).manual_seed(1)
bt = torch.randperm(nb, generator=torch.Generator().manual_seed(7))[: B * nblk].to(torch.int32).reshape(B, nblk).to(dev)
lens = torch.full((B,), L, dtype=torch.int32, device=dev)
caches = []
for _ in range(NL):
    ph = torch.randint(0, 0x78, (nb, 2, Hkv, bs, D), dtype=torch.uint8, device=dev, generator=g)
    caches.append(ph.permute(0, 1, 3, 2, 4).unbind(1))
q = torch.randn(B, Hq, D, device=dev, generator=g).to(torch.bfloat16)
out = torch.zeros(NL, B, Hq, D, dtype=torch.bfloat16, device=dev)
ks = torch.full((Hkv,), 0.01, device=dev); vs = torch.full((Hkv,), 0.01, device=dev)
nbytes = 2 * B * L * Hkv * D + 4 * B * Hq * D + B * nblk * 4
ref = None
VARIANTS = [(None, 2, True), (None, 2, False), (256, 2, True), (256, 2, False), (256, 8, True), (256, 4, True), (128, 4, True), (288, 2, True), (296, 2, True)]
if os.environ.get("B200_OVERSUB"):   # more CTAs than resident slots: the block scheduler balances the waves
    VARIANTS = [(None, 2, True), (256, 2, True), (384, 2, True), (384, 1, True), (512, 2, True), (512, 1, True), (512, 4, True), (768, 2, True), (768, 1, True), (1024, 2, True),
                (512, 2, False), (1024, 2, False)]
for ctas, cluster, fused in VARIANTS:
    lib.b200attn_set_option(b"cluster", cluster)
    ws = ops.DecodeWorkspace(dev, B, torch.uint8, num_ctas=ctas, group_size=Hq // Hkv)
    ws.make_plan(lens, Hq, Hkv, bt, bs)

    def body():
        for l in range(NL):
            if fused is True and ctas is None:
                ws.decode(q, caches[l][0], caches[l][1], bt, lens, out[l], ks, vs, D ** -0.5, fresh_tokens=1)     # ops.paged_decode's own routing
            else:
                ops.launch_split_decode(q, caches[l][0], caches[l][1], bt, lens, out[l], ws.plan, ws.workspace, ks, vs, D ** -0.5, ws.grid, None, None, 1, fused)
                if not fused:
                    ops.launch_merge(caches[l][0], lens, out[l], ws.plan, ws.workspace, vs, ws.grid)
    try:
        us, _ = _graph_time(body, NL)
        torch.cuda.synchronize()
        if ref is None:
            ref = out.clone()
        err = float((out.float() - ref.float()).abs().max())
        print(f"grid {ws.grid:4d} cluster {cluster} {'in-kernel merge' if fused else 'merge kernel   '}: {us:7.2f} us  {nbytes / us / 1e3:7.0f} GB/s  max|diff vs first| {err:.2e}", flush=True)
    except Exception as e:  # noqa: BLE001
        print(f"grid {ctas} cluster {cluster} fused {fused}: {type(e).__name__}: {str(e)[:200]}", flush=True)
lib.b200attn_set_option(b"cluster", 2)
