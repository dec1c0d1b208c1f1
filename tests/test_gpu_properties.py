"""GPU: size-independent properties at BASELINE.json's full sizes (where the numpy oracle would take minutes), plus a
sampled oracle check: the oracle is run on a few (sequence, head) pairs of the full-size problem only.

  * write -> gather round trip: the cache holds exactly quant(x/scale) of every token, at its slot, nowhere else;
  * split invariance: any decode grid (1 CTA ... 2 CTAs/SM) gives the same attention output up to fp32 re-association;
  * block-permutation invariance: relocating the physical blocks (and the block table with them) is bit-invariant;
  * convexity: a constant V plane returns that constant (softmax weights sum to one), for every head and sequence;
  * uniform attention: K = 0 gives the mean of V over the context;
  * linearity in V (16-bit KV): decode(V1 + V2) = decode(V1) + decode(V2) up to output rounding.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

D, BS = 128, 16


@pytest.fixture(scope="module")
def ops():
    assert torch.cuda.is_available()
    import ht_vllm_omni_b200.ops as ops

    ops.load_library()
    return ops


class Big:
    """C2-sized problem built on the device: B=64, L=4096 (or ragged), (Hq, Hkv) = (16, 2), fp8 or bf16 KV."""

    def __init__(self, ops, lens, Hq=16, Hkv=2, kv="fp8", seed=0):
        dev = "cuda"
        self.ops, self.Hq, self.Hkv, self.kv = ops, Hq, Hkv, kv
        self.lens = np.asarray(lens, np.int32)
        self.B = len(lens)
        g = torch.Generator(device=dev).manual_seed(4321 + seed)
        nblk = -(-int(self.lens.max()) // BS)
        need = int(sum(-(-int(x) // BS) for x in self.lens))
        self.nb = int(need * 1.1) + 2
        perm = torch.randperm(self.nb, generator=torch.Generator().manual_seed(99 + seed))
        bt = torch.zeros(self.B, nblk, dtype=torch.int32)
        o = 0
        for r, L in enumerate(self.lens):
            n = -(-int(L) // BS)
            bt[r, :n] = perm[o:o + n].to(torch.int32)
            o += n
        self.bt = bt.to(dev)
        self.d_lens = torch.from_numpy(self.lens).to(dev)
        self.T = int(self.lens.sum())
        self.key = torch.randn(self.T, Hkv, D, generator=g, device=dev).to(torch.bfloat16)
        self.value = torch.randn(self.T, Hkv, D, generator=g, device=dev).to(torch.bfloat16)
        self.q = torch.randn(self.B, Hq, D, generator=g, device=dev).to(torch.bfloat16)
        seq = torch.repeat_interleave(torch.arange(self.B), torch.from_numpy(self.lens).long())
        start = np.concatenate([[0], np.cumsum(self.lens)[:-1]])
        pos = torch.arange(self.T) - torch.from_numpy(start).long()[seq]
        self.slots = (bt.long()[seq, pos // BS] * BS + pos % BS).to(dev)
        self.seq, self.pos = seq, pos
        fp8 = kv == "fp8"
        self.dt = torch.uint8 if fp8 else torch.bfloat16
        self.ks = (self.key.float().abs().amax(dim=(0, 2)) / 448.0) if fp8 else torch.ones(1, device=dev)
        self.vs = (self.value.float().abs().amax(dim=(0, 2)) / 448.0) if fp8 else torch.ones(1, device=dev)
        self.phys = torch.zeros(self.nb, 2, Hkv, BS, D, dtype=self.dt, device=dev)
        self.kc, self.vc = self.phys.permute(0, 1, 3, 2, 4).unbind(1)
        self.sm_scale = D ** -0.5

    def clone_planes(self):
        """A second physical cache with the same strides (k/v planes must share one stride set)."""
        phys = self.phys.clone()
        return phys.permute(0, 1, 3, 2, 4).unbind(1)

    def write(self):
        torch.ops.b200attn.kv_write(self.key, self.value, self.kc, self.vc, self.slots, self.ks, self.vs)

    def decode(self, ctas=None, kc=None, vc=None, bt=None, q=None):
        ws = self.ops.DecodeWorkspace("cuda", self.B, self.dt, num_ctas=ctas)
        out = torch.full((self.B, self.Hq, D), float("nan"), dtype=torch.bfloat16, device="cuda")
        ws.make_plan(self.d_lens, self.Hq, self.Hkv, self.bt if bt is None else bt, BS)
        ws.decode(self.q if q is None else q, self.kc if kc is None else kc, self.vc if vc is None else vc,
                  self.bt if bt is None else bt, self.d_lens, out, self.ks, self.vs, self.sm_scale)
        torch.cuda.synchronize()
        return out


@pytest.fixture(scope="module")
def c2(ops):
    b = Big(ops, [4096] * 64, kv="fp8")
    b.write()
    torch.cuda.synchronize()
    return b


def test_full_size_write_round_trip(c2):
    """Every token's bytes sit at its slot and equal the e4m3 satfinite quantisation of x/scale (checked with torch's
    own cast on the device after clamping, which the oracle tests pin to be the same function in range)."""
    b = c2
    for plane, src, sc in ((b.kc, b.key, b.ks), (b.vc, b.value, b.vs)):
        want = (src.float() / sc[None, :, None]).clamp(-448, 448).to(torch.float8_e4m3fn).view(torch.uint8)
        got = plane[b.slots // BS, b.slots % BS]
        assert torch.equal(got, want)
    # blocks that no sequence owns stay zero
    owned = torch.zeros(b.nb, dtype=torch.bool, device="cuda")
    owned[(b.slots // BS).unique()] = True
    assert int(b.phys[~owned].abs().sum()) == 0


def test_full_size_sampled_oracle(c2):
    """Oracle on a sample of (sequence, kv head) pairs of the full-size problem."""
    from oracle import oracle as O

    b = c2
    out = b.decode().float().cpu().numpy()
    kc = b.kc.contiguous().cpu().numpy(); vc = b.vc.contiguous().cpu().numpy()
    bt = b.bt.cpu().numpy()
    rng = np.random.default_rng(0)
    for s in rng.choice(b.B, 3, replace=False):
        ref = O.paged_decode_exact(b.q[s:s + 1].float().cpu().numpy(), kc, vc, bt[s:s + 1], b.lens[s:s + 1], b.sm_scale,
                                   b.ks.cpu().numpy(), b.vs.cpu().numpy(), fp8=True)
        ulp = np.maximum(np.abs(ref[0]), 2.0 ** -14) * 2.0 ** -7
        assert np.all(np.abs(out[s] - ref[0]) <= 1e-2 + ulp), float(np.abs(out[s] - ref[0]).max())


def test_full_size_split_invariance(c2):
    base = c2.decode(ctas=None).float()
    assert not torch.isnan(base).any()
    for ctas in (1, 37, 148, 500):
        o = c2.decode(ctas=ctas).float()
        # identical maths, different association of the fp32 partial sums + one bf16 rounding
        assert float((o - base).abs().max()) <= 2.0 ** -9, ctas


def test_full_size_block_permutation_invariance(c2):
    b = c2
    base = b.decode()
    perm = torch.randperm(b.nb, device="cuda", generator=torch.Generator(device="cuda").manual_seed(3))
    inv = torch.empty_like(perm); inv[perm] = torch.arange(b.nb, device="cuda")
    phys2 = torch.empty_like(b.phys); phys2[perm] = b.phys       # block i moves to perm[i]
    kc2, vc2 = phys2.permute(0, 1, 3, 2, 4).unbind(1)
    bt2 = perm[b.bt.long()].to(torch.int32)
    o2 = b.decode(kc=kc2, vc=vc2, bt=bt2)
    assert torch.equal(o2, base)     # same plan, same arithmetic order: bit-identical


@pytest.mark.parametrize("kv", ["fp8", "bf16"])
def test_convexity_and_uniform_attention(ops, kv):
    lens = np.exp(np.random.default_rng(7).uniform(np.log(512), np.log(65536), 16)).astype(np.int32)   # ragged long tail
    b = Big(ops, lens, Hq=32, Hkv=4, kv=kv, seed=1)
    b.write()
    const = 0.75
    kc2, vconst = b.clone_planes()
    if kv == "fp8":
        vconst.fill_(int(torch.tensor(const).to(torch.float8_e4m3fn).view(torch.uint8).item()))
        expect = const * b.vs.repeat_interleave(b.Hq // b.Hkv)[None, :, None].expand(b.B, b.Hq, D)
    else:
        vconst.fill_(const)
        expect = torch.full((b.B, b.Hq, D), const, device="cuda")
    o = b.decode(kc=kc2, vc=vconst).float()
    assert float((o - expect.to(torch.bfloat16).float()).abs().max()) <= 2.0 ** -8 * float(expect.abs().max())
    # K = 0 -> uniform weights -> mean of V over the context
    kzero, vsame = b.clone_planes()
    kzero.zero_()
    o = b.decode(kc=kzero, vc=vsame).float()
    vsum = torch.zeros(b.B, b.Hkv, D, device="cuda", dtype=torch.float32)
    vdeq = (b.vc[b.slots // BS, b.slots % BS].view(torch.float8_e4m3fn).float() * b.vs[None, :, None]) if kv == "fp8" \
        else b.vc[b.slots // BS, b.slots % BS].float()
    vsum.index_add_(0, b.seq.to("cuda"), vdeq)
    mean = vsum / b.d_lens[:, None, None].float()
    mean = mean.repeat_interleave(b.Hq // b.Hkv, dim=1)
    assert float((o - mean).abs().max()) <= (1e-2 if kv == "fp8" else 2e-3)


def test_linearity_in_v_bf16(ops):
    b = Big(ops, [3000, 1111, 4096, 5], Hq=16, Hkv=2, kv="bf16", seed=2)
    b.write()
    k2, v2 = b.clone_planes()
    k3, v3 = b.clone_planes()
    v2.copy_((torch.randn(b.vc.shape, device="cuda", generator=torch.Generator(device="cuda").manual_seed(1)) * 0.5).to(torch.bfloat16))
    v3.copy_((b.vc.float() + v2.float()).to(torch.bfloat16))          # exact in fp32; rounded once
    o1, o2, o12 = b.decode().float(), b.decode(kc=k2, vc=v2).float(), b.decode(kc=k3, vc=v3).float()
    # bf16 rounding of V1+V2 (rel 2^-9 per element, averaged by the convex weights) + three output roundings
    assert float((o12 - (o1 + o2)).abs().max()) <= 2e-3 + 3 * 2.0 ** -8
