"""Time the kernels of the hot path in isolation on a B200 (CUDA events, L2-cold by rotating over > L2 of caches).

    python tests/tools/kernel_bench.py [--B 64 --L 4096 --Hq 16 --Hkv 2 --kv fp8 --layers 8 --ctas 0 --debug 0 --pdl 0]
Prints one line per configuration: avg us per launch and algorithmic GB/s for decode, merge, kv_write.
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import numpy as np
import torch

import ht_vllm_omni_b200.ops as ops
from bench import algorithmic_bytes


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--B", type=int, default=64); ap.add_argument("--L", type=int, default=4096)
    ap.add_argument("--Hq", type=int, default=16); ap.add_argument("--Hkv", type=int, default=2)
    ap.add_argument("--kv", default="fp8"); ap.add_argument("--layers", type=int, default=8)
    ap.add_argument("--ctas", default="0"); ap.add_argument("--debug", default="0"); ap.add_argument("--pdl", type=int, default=0)
    ap.add_argument("--tc", type=int, default=1); ap.add_argument("--graph", type=int, default=0); ap.add_argument("--reps", type=int, default=5); ap.add_argument("--ragged", type=int, default=0); ap.add_argument("--bs", type=int, default=16)
    a = ap.parse_args()
    dev = torch.device("cuda")
    lib = ops.load_library()
    B, L, Hq, Hkv, D, bs = a.B, a.L, a.Hq, a.Hkv, 128, a.bs
    fp8 = a.kv == "fp8"
    kv_t = torch.uint8 if fp8 else torch.bfloat16
    rng = np.random.default_rng(7)
    lens = np.full(B, L, np.int32) if not a.ragged else np.exp(rng.uniform(np.log(512), np.log(L), B)).astype(np.int32)
    nblk = -(-int(lens.max()) // bs)
    nb = int(B * nblk * 1.05) + 1
    bt = torch.randperm(nb, generator=torch.Generator().manual_seed(1))[: B * nblk].to(torch.int32).reshape(B, nblk).to(dev)
    d_lens = torch.from_numpy(lens).to(dev)
    caches = []
    for _ in range(a.layers):
        if fp8:
            phys = torch.randint(0, 0x78, (nb, 2, Hkv, bs, D), dtype=torch.uint8, device=dev)
        else:
            phys = torch.randn((nb, 2, Hkv, bs, D), dtype=torch.float32, device=dev).to(torch.bfloat16)
        caches.append(phys.permute(0, 1, 3, 2, 4).unbind(1))
    q = torch.randn(B, Hq, D, device=dev).to(torch.bfloat16)
    out = torch.zeros(B, Hq, D, dtype=torch.bfloat16, device=dev)
    ks = torch.full((Hkv,), 0.01, device=dev); vs = torch.full((Hkv,), 0.01, device=dev)
    kn = torch.randn(B, Hkv, D, device=dev).to(torch.bfloat16); vn = torch.randn(B, Hkv, D, device=dev).to(torch.bfloat16)
    slots = (bt.to(torch.int64)[torch.arange(B), torch.from_numpy((lens - 1) // bs).to(dev)] * bs + torch.from_numpy((lens - 1) % bs).to(dev)).contiguous()
    dec_b, wr_b = algorithmic_bytes(lens, Hq, Hkv, D, bs, 1 if fp8 else 2)
    lib.b200attn_set_option(b"pdl", a.pdl)
    lib.b200attn_set_option(b"tc", a.tc)
    for ctas in [int(x) for x in a.ctas.split(",")]:
        for dbg in [int(x) for x in a.debug.split(",")]:
            lib.b200attn_set_option(b"debug", dbg)
            ws = ops.DecodeWorkspace(dev, B, kv_t, num_ctas=ctas or None)
            ws.make_plan(d_lens, Hq, Hkv, bt, bs)
            torch.cuda.synchronize()

            def timed(fn):
                evs = []
                for r in range(a.reps + 1):
                    for l in range(a.layers):
                        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        e0.record(); fn(l); e1.record()
                        if r > 0:
                            evs.append((e0, e1))
                torch.cuda.synchronize()
                t = np.array([x.elapsed_time(y) for x, y in evs]) * 1e3
                return float(t.mean()), float(np.median(t)), float(t.min())

            td = timed(lambda l: ops.launch_split_decode(q, caches[l][0], caches[l][1], bt, d_lens, out, ws.plan, ws.workspace, ks, vs, D ** -0.5, ws.grid))
            tm = timed(lambda l: ops.launch_merge(caches[l][0], d_lens, out, ws.plan, ws.workspace, vs, ws.grid))
            tw = timed(lambda l: torch.ops.b200attn.kv_write(kn, vn, caches[l][0], caches[l][1], slots, ks, vs))
            tall = timed(lambda l: (torch.ops.b200attn.kv_write(kn, vn, caches[l][0], caches[l][1], slots, ks, vs), ws.decode(q, caches[l][0], caches[l][1], bt, d_lens, out, ks, vs, D ** -0.5)))
            if a.graph:
                def graph_time(body, iters=10):
                    side = torch.cuda.Stream(); side.wait_stream(torch.cuda.current_stream())
                    with torch.cuda.stream(side):
                        body()
                    torch.cuda.current_stream().wait_stream(side); torch.cuda.synchronize()
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g):
                        body()
                    for _ in range(3):
                        g.replay()
                    torch.cuda.synchronize()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    for _ in range(iters):
                        g.replay()
                    e1.record(); torch.cuda.synchronize()
                    return e0.elapsed_time(e1) * 1e3 / iters / a.layers
                dec = lambda l: ops.launch_split_decode(q, caches[l][0], caches[l][1], bt, d_lens, out, ws.plan, ws.workspace, ks, vs, D ** -0.5, ws.grid)
                mrg = lambda l: ops.launch_merge(caches[l][0], d_lens, out, ws.plan, ws.workspace, vs, ws.grid)
                wrt = lambda l: torch.ops.b200attn.kv_write(kn, vn, caches[l][0], caches[l][1], slots, ks, vs)
                for pdl in (0, 1):
                    lib.b200attn_set_option(b"pdl", pdl)
                    t_d = graph_time(lambda: [dec(l) for l in range(a.layers)])
                    t_dm = graph_time(lambda: [(dec(l), mrg(l)) for l in range(a.layers)])
                    t_wdm = graph_time(lambda: [(wrt(l), dec(l), mrg(l)) for l in range(a.layers)])
                    decf = lambda l, fr: ops.launch_split_decode(q, caches[l][0], caches[l][1], bt, d_lens, out, ws.plan, ws.workspace, ks, vs, D ** -0.5, ws.grid, None, None, fr, True)
                    t_f = graph_time(lambda: [decf(l, -1) for l in range(a.layers)])
                    t_wf = graph_time(lambda: [(wrt(l), decf(l, -1)) for l in range(a.layers)])
                    t_wf1 = graph_time(lambda: [(wrt(l), decf(l, 1)) for l in range(a.layers)])
                    print(f"  graph/layer pdl={pdl}: FUSED-merge decode {t_f:.2f} | write+fused {t_wf:.2f} | write+fused(fresh=1, early prefetch) {t_wf1:.2f} us => all {(dec_b + wr_b) / t_wf1 / 1e3:.0f} GB/s", flush=True)
                    t_w = graph_time(lambda: [wrt(l) for l in range(a.layers)])
                    t_m = graph_time(lambda: [mrg(l) for l in range(a.layers)])
                    print(f"  graph/layer pdl={pdl}: decode {t_d:.2f} | decode+merge {t_dm:.2f} | write+decode+merge {t_wdm:.2f} | write only {t_w:.2f} | merge only {t_m:.2f} us"
                          f"  => decode {dec_b / t_d / 1e3:.0f} GB/s, all {(dec_b + wr_b) / t_wdm / 1e3:.0f} GB/s", flush=True)
                lib.b200attn_set_option(b"pdl", a.pdl)
            print(f"B={B} L={L} ({Hq},{Hkv}) {a.kv} tc={a.tc} ctas={ws.grid} debug={dbg} pdl={a.pdl} | decode mean {td[0]:.2f} med {td[1]:.2f} min {td[2]:.2f} us "
                  f"= {dec_b / td[1] / 1e3:.0f} GB/s | merge {tm[1]:.2f} us | kv_write {tw[1]:.2f} us | write+decode+merge {tall[1]:.2f} us", flush=True)
    lib.b200attn_set_option(b"debug", 0)


if __name__ == "__main__":
    main()
