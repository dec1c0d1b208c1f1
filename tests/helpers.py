"""Shared builders for the GPU parity tests: seeded synthetic paged-KV problems in torch, with numpy
mirrors for the oracle.  The CUDA path is always driven through torch.ops.b200attn.* (i.e. through the C ABI)."""
from __future__ import annotations

import numpy as np
import torch

from oracle import oracle as O

ACT = {"bf16": torch.bfloat16, "fp16": torch.float16}


def to_np_f32(t: torch.Tensor) -> np.ndarray:
    return t.detach().float().cpu().numpy()


def alloc_cache(nb, bs, Hkv, D, kv, layout="HND", device="cuda", fill=0):
    """Logical [nb, 2, bs, Hkv, D] (vLLM shape) over HND or NHD physical memory; returns (kv_cache, k_view, v_view)."""
    dt = torch.uint8 if kv == "fp8" else ACT[kv]
    if layout == "HND":
        phys = torch.full((nb, 2, Hkv, bs, D), fill, dtype=dt, device=device)
        cache = phys.permute(0, 1, 3, 2, 4)
    else:
        cache = torch.full((nb, 2, bs, Hkv, D), fill, dtype=dt, device=device)
    k, v = cache.unbind(1)
    return cache, k, v


class Case:
    def __init__(self, lens, Hq, Hkv, D=128, bs=16, kv="fp8", act="bf16", seed=0, per_head=False, layout="HND",
                 device="cuda", dist="normal", spare=1.1):
        self.lens = np.asarray(lens, dtype=np.int32)
        self.B, self.Hq, self.Hkv, self.D, self.bs = len(lens), Hq, Hkv, D, bs
        self.kv, self.act, self.fp8 = kv, act, kv == "fp8"
        self.dist = dist
        self.device = device
        g = torch.Generator().manual_seed(1234 + seed)
        self.bt_np, self.nb = O.make_block_table(self.lens, bs, seed=99 + seed, spare=spare)
        T = int(self.lens.sum())
        adt = ACT[act]

        def rnd(*shape):
            if dist == "heavy":   # Student-t(3): exercises fp8 saturation
                chi2 = torch.randn(3, *shape, generator=g).square().sum(0)   # seeded (torch.distributions ignores `g`)
                x = torch.randn(*shape, generator=g) / torch.sqrt(chi2 / 3.0)
                return x.to(adt)
            return torch.randn(*shape, generator=g).to(adt)

        self.key = rnd(T, Hkv, D)
        self.value = rnd(T, Hkv, D)
        self.q = rnd(self.B, Hq, D)
        pos = [(r, p) for r in range(self.B) for p in range(int(self.lens[r]))]
        self.slots_np = O.slots_for_positions(self.bt_np, pos, bs)
        if self.fp8:
            if per_head:
                ks = self.key.float().abs().amax(dim=(0, 2)) / 448.0
                vs = self.value.float().abs().amax(dim=(0, 2)) / 448.0
            else:
                ks = (self.key.float().abs().max() / 448.0).reshape(1)
                vs = (self.value.float().abs().max() / 448.0).reshape(1)
            if dist == "heavy":     # make ~0.1% of the values saturate
                ks, vs = ks / 8, vs / 8
            self.k_scale, self.v_scale = ks.float().clamp_min(1e-6), vs.float().clamp_min(1e-6)
        else:
            self.k_scale = self.v_scale = torch.ones(1)
        self.sm_scale = float(D) ** -0.5
        self.layout = layout

    # ---- oracle side (numpy) ----
    def oracle_cache(self):
        shape = (self.nb, self.bs, self.Hkv, self.D)
        if self.fp8:
            kc, vc = np.zeros(shape, np.uint8), np.zeros(shape, np.uint8)
        else:
            kc, vc = np.zeros(shape, np.float32), np.zeros(shape, np.float32)
        O.kv_write_ref(to_np_f32(self.key), to_np_f32(self.value), kc, vc, self.slots_np,
                       self.k_scale.numpy(), self.v_scale.numpy(), fp8=self.fp8)
        return kc, vc

    def oracle_decode(self, kc, vc, style="ref"):
        args = (to_np_f32(self.q), kc, vc, self.bt_np, self.lens, self.sm_scale, self.k_scale.numpy(), self.v_scale.numpy())
        if style == "exact":
            return O.paged_decode_exact(*args, fp8=self.fp8)
        return O.paged_decode_ref(*args, fp8=self.fp8, act=self.act)

    # ---- CUDA side ----
    def cuda(self):
        d = self.device
        self.d_key, self.d_value, self.d_q = self.key.to(d), self.value.to(d), self.q.to(d)
        self.d_slots = torch.from_numpy(self.slots_np).to(d)
        self.d_bt = torch.from_numpy(self.bt_np).to(d)
        self.d_lens = torch.from_numpy(self.lens).to(d)
        self.d_ks, self.d_vs = self.k_scale.to(d), self.v_scale.to(d)
        self.cache, self.kc, self.vc = alloc_cache(self.nb, self.bs, self.Hkv, self.D, self.kv, self.layout, d)
        return self

    def cache_to_np(self, plane: torch.Tensor) -> np.ndarray:
        t = plane.contiguous().cpu()
        return t.numpy() if self.fp8 else t.float().numpy()

    def load_cache_from_np(self, kc, vc):
        """Put oracle-written cache contents on the device (decoupling decode tests from the write kernel)."""
        dt = torch.uint8 if self.fp8 else ACT[self.kv]
        self.kc.copy_(torch.from_numpy(kc).to(dt))
        self.vc.copy_(torch.from_numpy(vc).to(dt))
