"""Fused rope_kv_write vs the reference's two compiled ops (rotary_embedding + reshape_and_cache_flash) on one B200:
CUDA-graph replays of 20 layer-launches over distinct caches, CUDA events, us per layer.

    python tests/tools/rope_bench.py [--T 64,8192] [--Hq 16 --Hkv 2]
"""
import argparse, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import torch
import ht_vllm_omni_b200.ops  # noqa: F401
from oracle import oracle as O

ap = argparse.ArgumentParser()
ap.add_argument("--T", default="64,1024,8192"); ap.add_argument("--Hq", type=int, default=16); ap.add_argument("--Hkv", type=int, default=2)
a = ap.parse_args()
try:
    import vllm._custom_ops as cops
except Exception as e:  # noqa: BLE001
    cops = None
    print("vLLM ops unavailable:", e)
D, bs, L = 128, 16, 20
Hq, Hkv = a.Hq, a.Hkv
cache = torch.from_numpy(O.rope_cos_sin_cache(D, 32768, 1e6, "bf16")).bfloat16().cuda()
ks = torch.full((1,), 0.03, device="cuda"); vs = torch.full((1,), 0.02, device="cuda")


def graph_us(body, iters=20):
    s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        body()
    torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
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
    return e0.elapsed_time(e1) * 1e3 / iters / L


print("| T | Hq:Hkv | fused rope_kv_write us | kv_write alone us | vLLM rotary_embedding + reshape_and_cache_flash us | algorithmic GB/s (fused) |")
print("|---|---|---|---|---|---|")
for T in [int(x) for x in a.T.split(",")]:
    nb = T // bs + 2
    qkv = [torch.randn(T, (Hq + 2 * Hkv) * D, device="cuda").bfloat16() for _ in range(L)]
    caches = [torch.zeros((nb, 2, bs, Hkv, D), dtype=torch.uint8, device="cuda") for _ in range(L)]
    pos = torch.randint(0, 32768, (T,), device="cuda")
    slots = torch.randperm(nb * bs, device="cuda")[:T].to(torch.int64)

    def split(x):
        q, k, v = x.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
        return q, k, v

    def fused():
        for l in range(L):
            q, k, v = split(qkv[l]); kc, vc = caches[l].unbind(1)
            torch.ops.b200attn.rope_kv_write(q.view(T, Hq, D), k.view(T, Hkv, D), v.view(T, Hkv, D), pos, cache, True, kc, vc, slots, ks, vs)

    def write_only():
        for l in range(L):
            q, k, v = split(qkv[l]); kc, vc = caches[l].unbind(1)
            torch.ops.b200attn.kv_write(k.view(T, Hkv, D), v.view(T, Hkv, D), kc, vc, slots, ks, vs)

    def ref():
        for l in range(L):
            q, k, v = split(qkv[l]); kc, vc = caches[l].unbind(1)
            cops.rotary_embedding(pos, q, k, D, cache, True)
            cops.reshape_and_cache_flash(k.view(T, Hkv, D), v.view(T, Hkv, D), kc.view(torch.float8_e4m3fn), vc.view(torch.float8_e4m3fn), slots, "fp8", ks, vs)

    tf, tw = graph_us(fused), graph_us(write_only)
    tr = graph_us(ref) if cops is not None else float("nan")
    byts = T * (Hq + Hkv) * D * 2 * 2 + T * Hkv * D * (2 + 2) + T * D * 2 + 16 * T
    print(f"| {T} | {Hq}:{Hkv} | {tf:.2f} | {tw:.2f} | {tr:.2f} | {byts / tf / 1e3:.0f} |", flush=True)
