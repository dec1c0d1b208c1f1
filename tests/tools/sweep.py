"""BASELINE configs[2] on one B200: the synthetic paged-decode sweep (batch x context, block 16, GQA 8:1, D 128, bf16 vs FP8
KV).  Each point times the decode op (split-KV kernel + merge, as ops.paged_decode launches them) as a CUDA graph of `layers` launches over distinct caches (working set
>= 2x L2 so nothing is reused), CUDA events around replays, and reports algorithmic GB/s (bench.algorithmic_bytes).

    python tests/tools/sweep.py [--points "1x1024,64x4096,..."] [--kv fp8,bf16] [--Hq 16 --Hkv 2] [--ctas 0] [--tc 1]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import numpy as np
import torch

import ht_vllm_omni_b200.ops as ops
from bench import algorithmic_bytes, load_peaks

DEFAULT = "1x1024,1x16384,1x131072,8x4096,8x32768,64x1024,64x4096,64x16384,256x1024,256x4096,16x65536,4x131072"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--points", default=DEFAULT)
    ap.add_argument("--kv", default="fp8,bf16")
    ap.add_argument("--Hq", type=int, default=16); ap.add_argument("--Hkv", type=int, default=2)
    ap.add_argument("--ctas", type=int, default=0); ap.add_argument("--tc", type=int, default=1)
    ap.add_argument("--ragged", type=int, default=0, help="1: log-uniform lengths in [512, L] (configs[4])")
    a = ap.parse_args()
    dev = torch.device("cuda")
    lib = ops.load_library()
    lib.b200attn_set_option(b"tc", a.tc)
    lib.b200attn_set_option(b"pdl", 1)
    peak, _ = load_peaks()
    Hq, Hkv, D, bs = a.Hq, a.Hkv, 128, 16
    print(f"| B | L | KV | layers | us/launch | GB/s | % of measured HBM ({peak:.0f} GB/s) |")
    print("|---|---|---|---|---|---|---|")
    for kv in a.kv.split(","):
        fp8 = kv == "fp8"
        s_kv = 1 if fp8 else 2
        kv_t = torch.uint8 if fp8 else torch.bfloat16
        for pt in a.points.split(","):
            B, L = (int(x) for x in pt.split("x"))
            rng = np.random.default_rng(7)
            lens = np.full(B, L, np.int32) if not a.ragged else np.exp(rng.uniform(np.log(512), np.log(L), B)).astype(np.int32)
            nblk = -(-int(lens.max()) // bs)
            nb = B * nblk + 8
            layer_bytes = nb * 2 * Hkv * bs * D * s_kv
            layers = int(min(16, max(2, -(-300e6 // layer_bytes))))
            if layers * layer_bytes > 40e9:
                layers = max(1, int(40e9 // layer_bytes))
            bt = torch.randperm(nb, generator=torch.Generator().manual_seed(1))[: B * nblk].to(torch.int32).reshape(B, nblk).to(dev)
            d_lens = torch.from_numpy(lens).to(dev)
            caches = []
            for _ in range(layers):
                if fp8:
                    phys = torch.randint(0, 0x78, (nb, 2, Hkv, bs, D), dtype=torch.uint8, device=dev)
                else:
                    phys = torch.empty((nb, 2, Hkv, bs, D), dtype=torch.bfloat16, device=dev).normal_()
                caches.append(phys.permute(0, 1, 3, 2, 4).unbind(1))
            q = torch.randn(B, Hq, D, device=dev).to(torch.bfloat16)
            out = torch.zeros(B, Hq, D, dtype=torch.bfloat16, device=dev)
            ks = torch.full((Hkv,), 0.01, device=dev); vs = torch.full((Hkv,), 0.01, device=dev)
            ws = ops.DecodeWorkspace(dev, B, kv_t, num_ctas=a.ctas or None, group_size=Hq // Hkv)
            ws.make_plan(d_lens, Hq, Hkv, bt, bs)

            def body():
                for l in range(layers):
                    ws.decode(q, caches[l][0], caches[l][1], bt, d_lens, out, ks, vs, D ** -0.5)   # public op (merge policy included)
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
            iters = 10
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(iters):
                g.replay()
            e1.record(); torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / iters / layers
            dec_b, _ = algorithmic_bytes(lens, Hq, Hkv, D, bs, s_kv)
            gbs = dec_b / us / 1e3
            tag = f"{L}" + (" (ragged 512..L)" if a.ragged else "")
            print(f"| {B} | {tag} | {kv} | {layers} | {us:.2f} | {gbs:.0f} | {100 * gbs / peak:.1f} |", flush=True)
            del caches, g
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
