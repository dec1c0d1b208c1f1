"""KV hand-off: kv_handoff.gather_request_kv (one launch + one pinned D2H) vs the reference's per-layer loop
(index -> flatten -> [:seq_len] -> .cpu().contiguous(), kv_transfer_manager.py:262-287 restated in torch) on one B200.
Wall-clock per request (the call is synchronous in the reference too), median of 10.

    python tests/tools/handoff_bench.py [--L 20 --ctx 4096,32768 --Hkv 2 --kv fp8]
"""
import argparse, os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import numpy as np
import torch
from ht_vllm_omni_b200 import kv_handoff

ap = argparse.ArgumentParser()
ap.add_argument("--L", type=int, default=20); ap.add_argument("--ctx", default="4096,32768"); ap.add_argument("--Hkv", type=int, default=2)
ap.add_argument("--kv", default="fp8")
a = ap.parse_args()
bs, D = 16, 128
dt = torch.uint8 if a.kv == "fp8" else torch.bfloat16
print("| layers | context | KV | bytes/request | reference loop ms | gather ms | speed-up | gather GB/s (incl. D2H) |")
print("|---|---|---|---|---|---|---|---|")
for ctx in [int(x) for x in a.ctx.split(",")]:
    nblk = ctx // bs
    nb = nblk * 4
    caches = [torch.zeros((nb, 2, a.Hkv, bs, D), dtype=dt, device="cuda").permute(0, 1, 3, 2, 4) for _ in range(a.L)]
    for c in caches:
        c.view(torch.uint8).random_(0, 255) if dt == torch.uint8 else c.normal_()
    ids = torch.randperm(nb)[:nblk].tolist()
    seq_len = ctx - 5

    def ref():
        ks, vs = [], []
        for c in caches:
            kb, vb = c[:, 0], c[:, 1]
            fk, fv = kb[ids].flatten(0, 1)[:seq_len], vb[ids].flatten(0, 1)[:seq_len]
            ks.append(fk.detach().cpu().contiguous()); vs.append(fv.detach().cpu().contiguous())
        return ks, vs

    def ours():
        return kv_handoff.gather_request_kv(ids, seq_len, caches)

    r, o = ref(), ours()
    assert all(torch.equal(x.view(torch.uint8), y.view(torch.uint8)) for x, y in zip(r[0] + r[1], o[0] + o[1]))

    def med(fn):
        ts = []
        for _ in range(10):
            torch.cuda.synchronize(); t0 = time.perf_counter(); fn(); torch.cuda.synchronize(); ts.append(time.perf_counter() - t0)
        return float(np.median(ts))
    tr, to = med(ref), med(ours)
    byts = a.L * 2 * seq_len * a.Hkv * D * caches[0].element_size()
    print(f"| {a.L} | {ctx} | {a.kv} | {byts / 1e6:.1f} MB | {tr * 1e3:.2f} | {to * 1e3:.2f} | {tr / to:.1f}x | {byts / to / 1e9:.1f} |", flush=True)
