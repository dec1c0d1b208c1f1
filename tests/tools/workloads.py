"""Synthetic paged-KV workloads shared by bench.py and the profiling tools: seeded, HND physical caches,
random block permutation, N(0,1) tokens written THROUGH the KV-write path (fill="write") or random NaN-free e4m3 bytes /
N(0,1) bf16 (fill="random", fast; for timing-only tools)."""
from __future__ import annotations

import numpy as np
import torch

WORKLOADS = {
    # name: B, L (int or "ragged:<lo>:<hi>:<seed>" / "tail"), Hq, Hkv, D, block_size, kv, layers
    "c2": dict(B=64, L=4096, Hq=16, Hkv=2, D=128, bs=16, kv="fp8", layers=20,
               desc="Qwen3-TTS talker streaming decode, FP8 KV, batch 64, 4k context (BASELINE configs[1])"),
    "c2-bf16": dict(B=64, L=4096, Hq=16, Hkv=2, D=128, bs=16, kv="bf16", layers=20, desc="C2 with bf16 KV"),
    "thinker": dict(B=32, L=32768, Hq=32, Hkv=4, D=128, bs=16, kv="fp8", layers=2,
                    desc="Qwen3-Omni thinker decode, FP8 KV, 32k context (BASELINE configs[3], 2 layers resident)"),
    "c1": dict(B=1, L=2048, Hq=16, Hkv=2, D=128, bs=16, kv="bf16", layers=20, desc="talker B1 2k bf16 (configs[0])"),
    "b1-128k": dict(B=1, L=131072, Hq=32, Hkv=4, D=128, bs=16, kv="fp8", layers=8, desc="B 1 x 128k fp8 (configs[2] corner)"),
    "ragged": dict(B=64, L="ragged:512:65536:7", Hq=32, Hkv=4, D=128, bs=16, kv="fp8", layers=2,
                   desc="thinker, ragged log-uniform 512-64k seed 7 (configs[4])"),
    "tail": dict(B=64, L="tail", Hq=32, Hkv=4, D=128, bs=16, kv="fp8", layers=8, desc="1 x 64k + 63 x 512 (configs[4] adversarial)"),
}


def seq_lens_of(w) -> np.ndarray:
    B, L = w["B"], w["L"]
    if isinstance(L, int):
        return np.full(B, L, dtype=np.int32)
    if L == "tail":
        a = np.full(B, 512, dtype=np.int32)
        a[0] = 65536
        return a
    _, lo, hi, seed = L.split(":")
    rng = np.random.default_rng(int(seed))
    return np.exp(rng.uniform(np.log(int(lo)), np.log(int(hi)), B)).astype(np.int32)


def algorithmic_bytes(lens, Hq, Hkv, D, bs, s_kv):
    """BASELINE.md 'Algorithmic bytes': (decode bytes, kv-write bytes) per layer call."""
    lens = np.asarray(lens, dtype=np.int64)
    B = len(lens)
    dec = int((2 * lens * Hkv * D * s_kv).sum() + 2 * B * Hq * D * 2 + (-(-lens // bs) * 4).sum() + 4 * B)
    wr = int(B * Hkv * D * (2 * 2 + 2 * s_kv) + 8 * B)
    return dec, wr


class Problem:
    """Device tensors of one workload on one device."""

    def __init__(self, w, dev, seed=1234, fill="write", layers=None):
        import ht_vllm_omni_b200.ops  # noqa: F401  (registers torch.ops.b200attn.*)

        self.w = w
        B, Hq, Hkv, D, bs = w["B"], w["Hq"], w["Hkv"], w["D"], w["bs"]
        NL = layers or w["layers"]
        self.NL = NL
        fp8 = w["kv"] == "fp8"
        self.fp8 = fp8
        self.kv_t = torch.uint8 if fp8 else torch.bfloat16
        self.lens_np = seq_lens_of(w)
        lens = self.lens_np
        g = torch.Generator(device=dev).manual_seed(seed)
        nblk = -(-lens.astype(np.int64) // bs)
        max_blk = int(nblk.max())
        total_blk = int(nblk.sum())
        nb = int(total_blk * 1.1) + 1
        perm = torch.randperm(nb, generator=torch.Generator().manual_seed(99))[:total_blk].to(torch.int32)
        bt = torch.zeros(B, max_blk, dtype=torch.int32)
        off = 0
        for s in range(B):
            bt[s, : nblk[s]] = perm[off: off + nblk[s]]
            off += int(nblk[s])
        self.bt = bt.to(dev)
        self.lens = torch.from_numpy(lens).to(dev)
        self.k_scale = torch.full((Hkv,), 4.5 / 448.0, dtype=torch.float32, device=dev)   # N(0,1) tokens: |x| <= 4.5 sigma
        self.v_scale = torch.full((Hkv,), 4.5 / 448.0, dtype=torch.float32, device=dev)
        self.phys, self.planes = [], []
        for _ in range(NL):
            if fill == "random":
                if fp8:
                    phys = torch.randint(0, 0x78, (nb, 2, Hkv, bs, D), dtype=torch.uint8, device=dev, generator=g)
                    phys |= (torch.randint(0, 2, (nb, 2, Hkv, bs, D), dtype=torch.uint8, device=dev, generator=g) << 7)
                else:
                    phys = torch.randn((nb, 2, Hkv, bs, D), dtype=torch.float32, device=dev, generator=g).to(torch.bfloat16)
                kc, vc = phys.permute(0, 1, 3, 2, 4).unbind(1)
            else:
                phys = torch.zeros((nb, 2, Hkv, bs, D), dtype=self.kv_t, device=dev)
                kc, vc = phys.permute(0, 1, 3, 2, 4).unbind(1)
                for s in range(B):
                    L = int(lens[s])
                    pos = torch.arange(L, dtype=torch.int64)
                    slots = (bt[s].to(torch.int64)[pos // bs] * bs + pos % bs).to(dev)
                    chunk = 1 << 16
                    for s0 in range(0, L, chunk):
                        n = min(chunk, L - s0)
                        kk = torch.randn(n, Hkv, D, generator=g, device=dev, dtype=torch.float32).to(torch.bfloat16)
                        vv = torch.randn(n, Hkv, D, generator=g, device=dev, dtype=torch.float32).to(torch.bfloat16)
                        torch.ops.b200attn.kv_write(kk, vv, kc, vc, slots[s0:s0 + n], self.k_scale, self.v_scale)
            self.phys.append(phys)
            self.planes.append((kc, vc))
        torch.cuda.synchronize(dev)
        self.q = torch.randn(NL, B, Hq, D, generator=g, device=dev, dtype=torch.float32).to(torch.bfloat16)
        self.kn = torch.randn(NL, B, Hkv, D, generator=g, device=dev, dtype=torch.float32).to(torch.bfloat16)
        self.vn = torch.randn(NL, B, Hkv, D, generator=g, device=dev, dtype=torch.float32).to(torch.bfloat16)
        self.out = torch.zeros(NL, B, Hq, D, dtype=torch.bfloat16, device=dev)
        last = torch.from_numpy(lens.astype(np.int64) - 1)
        self.step_slots = (bt.to(torch.int64)[torch.arange(B), last // bs] * bs + last % bs).to(dev)   # newest token per sequence
        self.sm_scale = D ** -0.5
        self.dec_bytes, self.wr_bytes = algorithmic_bytes(lens, Hq, Hkv, D, bs, 1 if fp8 else 2)
