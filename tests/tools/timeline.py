"""Debug: per-CTA %globaltimer timeline of the decode kernel (entry, plan read, first TMA, first data, last tile, exit)."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import ctypes, numpy as np, torch
import ht_vllm_omni_b200.ops as ops
lib = ops.load_library()
dev = torch.device("cuda")
B, L, Hq, Hkv, D, bs = 64, 4096, 16, 2, 128, 16
nblk = L // bs; nb = B * nblk + 8
layers = 6
bt = torch.randperm(nb, generator=torch.Generator().manual_seed(1))[: B * nblk].to(torch.int32).reshape(B, nblk).to(dev)
lens = torch.full((B,), L, dtype=torch.int32, device=dev)
caches = [torch.randint(0, 0x78, (nb, 2, Hkv, bs, D), dtype=torch.uint8, device=dev).permute(0, 1, 3, 2, 4).unbind(1) for _ in range(layers)]
q = torch.randn(B, Hq, D, device=dev).to(torch.bfloat16); out = torch.zeros_like(q)
ks = torch.full((Hkv,), 0.01, device=dev); vs = torch.full((Hkv,), 0.01, device=dev)
ws = ops.DecodeWorkspace(dev, B, torch.uint8, num_ctas=int(os.environ.get("B200_CTAS", "0")) or None); ws.make_plan(lens, Hq, Hkv, bt, bs)
buf = torch.zeros(ws.grid * 8, dtype=torch.int64, device=dev)
lib.b200attn_set_option(b"pdl", 0)
lib.b200attn_set_option(b"tc", int(os.environ.get("B200_TC", "1")))
fn = lib.b200attn_debug_set_timeline_buffer; fn.argtypes = [ctypes.c_void_p]; fn.restype = None
for dbg in (0,):
    lib.b200attn_set_option(b"debug", dbg)
    for l in range(layers):
        fn(buf.data_ptr() if l == layers - 1 else None)
        ops.launch_split_decode(q, caches[l][0], caches[l][1], bt, lens, out, ws.plan, ws.workspace, ks, vs, D ** -0.5, ws.grid)
    torch.cuda.synchronize()
    t = buf.cpu().numpy().reshape(-1, 8).astype(np.float64)
    act = t[:, 1] > 0
    t0 = t[act, 0].min()
    names = ["entry", "plan read", "first TMA issued", "first data landed", "last tile done", "exit", "seq_len known", "block ids known"]
    print(f"debug={dbg} active CTAs {act.sum()}")
    print(f"  softmax warp 0: cycles waiting for S (total) p50 {np.median(t[act,6]):.0f}, cycles of work per tile p50 {np.median(t[act,7]):.0f} max {t[act,7].max():.0f}")
    for i, n in enumerate(names[:6]):
        v = (t[act, i] - t0) / 1e3
        print(f"  {n:18s} min {v.min():7.2f}  p50 {np.median(v):7.2f}  p90 {np.percentile(v,90):7.2f}  max {v.max():7.2f} us")
fn(None)
