"""Debug: %globaltimer timelines of CONSECUTIVE decode launches on one stream (PDL on), to see how the prologue / tail of
neighbouring kernels overlap.  Env: B200_TC, B200_FUSED (1), B200_FRESH (-1), B200_WRITE (0: decode only, 1: kv_write between)."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import ctypes, numpy as np, torch
import ht_vllm_omni_b200.ops as ops
lib = ops.load_library()
dev = torch.device("cuda")
B, L, Hq, Hkv, D, bs = 64, 4096, 16, 2, 128, 16
nblk = L // bs; nb = B * nblk + 8
layers = 8
fused = int(os.environ.get("B200_FUSED", "1")); fresh = int(os.environ.get("B200_FRESH", "-1")); wr = int(os.environ.get("B200_WRITE", "0"))
bt = torch.randperm(nb, generator=torch.Generator().manual_seed(1))[: B * nblk].to(torch.int32).reshape(B, nblk).to(dev)
lens = torch.full((B,), L, dtype=torch.int32, device=dev)
caches = [torch.randint(0, 0x78, (nb, 2, Hkv, bs, D), dtype=torch.uint8, device=dev).permute(0, 1, 3, 2, 4).unbind(1) for _ in range(layers)]
q = torch.randn(B, Hq, D, device=dev).to(torch.bfloat16); out = torch.zeros_like(q)
kn = torch.randn(B, Hkv, D, device=dev).to(torch.bfloat16); vn = torch.randn(B, Hkv, D, device=dev).to(torch.bfloat16)
slots = (bt.to(torch.int64)[:, (L - 1) // bs] * bs + (L - 1) % bs).contiguous()
ks = torch.full((Hkv,), 0.01, device=dev); vs = torch.full((Hkv,), 0.01, device=dev)
ws = ops.DecodeWorkspace(dev, B, torch.uint8, num_ctas=int(os.environ.get("B200_CTAS", "0")) or None); ws.make_plan(lens, Hq, Hkv, bt, bs)
nc = ws.grid     # the grid the plan was cut for (aligned_grid), which is what the launch must use
buf = torch.zeros(layers, nc * 16, dtype=torch.int64, device=dev)
lib.b200attn_set_option(b"pdl", int(os.environ.get("B200_PDL", "1")))
lib.b200attn_set_option(b"tc", int(os.environ.get("B200_TC", "1")))
lib.b200attn_set_option(b"cluster", int(os.environ.get("B200_CLUSTER", "2")))
fn = lib.b200attn_debug_set_timeline_buffer; fn.argtypes = [ctypes.c_void_p]; fn.restype = None
def chain():
    for l in range(layers):
        fn(buf[l].data_ptr())
        if wr:
            torch.ops.b200attn.kv_write(kn, vn, caches[l][0], caches[l][1], slots, ks, vs)
        ops.launch_split_decode(q, caches[l][0], caches[l][1], bt, lens, out, ws.plan, ws.workspace, ks, vs, D ** -0.5, nc, None, None, fresh, bool(fused))
if int(os.environ.get("B200_GRAPH", "1")):
    # the launches are captured (debug buffer pointers are baked into the kernel parameters) and replayed, so the gaps are
    # those of a CUDA graph, not of eager host launches
    side = torch.cuda.Stream(); side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        chain()
    torch.cuda.current_stream().wait_stream(side); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        chain()
    for rep in range(3):
        buf.zero_(); g.replay(); torch.cuda.synchronize()
else:
    for rep in range(2):
        buf.zero_(); chain(); torch.cuda.synchronize()
fn(None)
t = buf.cpu().numpy().reshape(layers, nc, 16).astype(np.float64)
t0 = t[0][t[0][:, 1] > 0][:, 0].min()
print(f"tc={os.environ.get('B200_TC','1')} fused={fused} fresh={fresh} write={wr} ctas={nc}: per launch, us relative to launch 0's first CTA entry")
prev_exit = None
for l in range(layers):
    a = t[l][t[l][:, 1] > 0]
    e = (a[:, 0] - t0) / 1e3; d = (a[:, 3] - t0) / 1e3; lt = (a[:, 4] - t0) / 1e3; x = (a[:, 5] - t0) / 1e3
    w6 = a[:, 6] / 1e3; w7 = a[:, 7] / 1e3
    if not (w7 > 0).any(): w7 = w7 + 1e-9
    slow = np.argsort(lt)[-3:]
    print(f"            publish calls per CTA: p50 {np.median(w6):5.2f} p90 {np.percentile(w6,90):5.2f} max {w6.max():5.2f} us | merge calls: p50 {np.median(w7[w7>0]):5.2f} p90 {np.percentile(w7[w7>0],90):5.2f} max {w7.max():5.2f} us | slowest: " + ", ".join(f"lt {lt[i]:.1f} pub {w6[i]:.2f} mrg {w7[i]:.2f}" for i in slow))
    gap = "" if prev_exit is None else f" | period {x.max() - prev_exit:6.2f}"
    rel = lambda k: (a[:, k] - a[:, 0]) / 1e3          # us from the CTA's own entry
    def q(v, f=np.median):
        v = v[(v > 0) & (v < 1e6)]
        return float(f(v)) if len(v) else float("nan")
    print(f"  launch {l}: entry min {e.min():7.2f} p50 {np.median(e):7.2f} max {e.max():7.2f} | from CTA entry (p50): set-up {q(rel(1)):5.2f}, plan entry {q(rel(8)):5.2f}, ring slot free {q(rel(14)):5.2f}, K TMA issued {q(rel(15)):5.2f}, V TMA issued {q(rel(2)):5.2f}, "
          f"Q ready {q(rel(9)):5.2f}, first QK issued {q(rel(10)):5.2f}, first S tile {q(rel(3)):5.2f} (max {q(rel(3), np.max):5.2f}) | first data p50 {np.median(d):7.2f} | "
          f"last tile p50 {np.median(lt):7.2f} max {lt.max():7.2f} | tail from last tile (p50/max): O read {q(rel(11) - rel(4)):4.2f}/{q(rel(11) - rel(4), np.max):4.2f}, "
          f"pushed {q(rel(13) - rel(4)):4.2f}/{q(rel(13) - rel(4), np.max):4.2f}, received {q(rel(12) - rel(4)):4.2f}/{q(rel(12) - rel(4), np.max):4.2f}, exit {q(rel(5) - rel(4)):4.2f}/{q(rel(5) - rel(4), np.max):4.2f} | "
          f"exit min {x.min():7.2f} max {x.max():7.2f}{gap}")
    prev_exit = x.max()
    if int(os.environ.get("B200_CLUSTER", "2")) == 2 and nc % 2 == 0 and len(a) == nc:
        # what dynamic tile sharing inside a cluster could give: both halves of a pair would finish at the mean of their
        # last-tile times instead of the later one
        pa, pb = lt[0::2], lt[1::2]
        cur, bal = np.maximum(pa, pb), (pa + pb) / 2
        print(f"            pair view: later half - earlier half p50 {np.median(np.abs(pa - pb)):.2f} max {np.abs(pa - pb).max():.2f} us; "
              f"slowest pair now {cur.max():.2f}, if balanced {bal.max():.2f} (p50 {np.median(cur):.2f} -> {np.median(bal):.2f}); "
              f"lower half later in {int((pa > pb).sum())} of {len(pa)} pairs")
    if l == layers - 1 and int(os.environ.get("B200_CLUSTER", "2")) > 1:
        # per-SM view: do CTAs that have their SM to themselves finish earlier?
        smid = a[:, 6].astype(np.int64)
        cnt = np.bincount(smid, minlength=160)
        ltr = rel(4)
        for k in (1, 2):
            sel = cnt[smid] == k
            if sel.any():
                print(f"            CTAs on SMs holding {k} CTA(s): n={int(sel.sum())}, first S tile p50 {np.median(rel(3)[sel]):5.2f}, last tile (from entry) p10 {np.percentile(ltr[sel],10):5.2f} p50 {np.median(ltr[sel]):5.2f} p90 {np.percentile(ltr[sel],90):5.2f} max {ltr[sel].max():5.2f}")
        print("            last tile (from entry) histogram, 0.5 us bins from", f"{ltr.min():.2f}:", np.histogram(ltr, bins=np.arange(ltr.min(), ltr.max() + 0.5, 0.5))[0].tolist())
