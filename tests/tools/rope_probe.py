"""Which arithmetic does vLLM's compiled rotary_embedding op use?  Compares it with candidate roundings (run on the GPU box)."""
import torch, sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import vllm._custom_ops as cops
from oracle import oracle as O
for dt in (torch.bfloat16, torch.float16):
    for neox in (True, False):
        T, H, D = 512, 8, 128
        g = torch.Generator().manual_seed(1)
        q = torch.randn(T, H * D, generator=g).to(dt).cuda()
        act = "bf16" if dt == torch.bfloat16 else "fp16"
        cache = torch.from_numpy(O.rope_cos_sin_cache(D, 8192, 1e6, act)).to(dt).cuda()
        pos = torch.randint(0, 8192, (T,), generator=g).cuda()
        ref = q.clone(); cops.rotary_embedding(pos, ref, None, D, cache, neox)
        x = q.view(T, H, D)
        cs = cache[pos]; c, s = cs[:, None, :D // 2], cs[:, None, D // 2:]
        x1, x2 = (x[..., :D // 2], x[..., D // 2:]) if neox else (x[..., 0::2], x[..., 1::2])
        def asm(o1, o2):
            return (torch.cat((o1, o2), -1) if neox else torch.stack((o1, o2), -1).flatten(-2)).reshape(T, -1)
        A = asm(x1 * c - x2 * s, x2 * c + x1 * s)
        f = lambda t: t.float()
        B = asm((f(x1) * f(c) - f(x2) * f(s)).to(dt), (f(x2) * f(c) + f(x1) * f(s)).to(dt))
        d = lambda t: t.double()
        Dd = asm((d(x1) * d(c) - d(x2) * d(s)).to(dt), (d(x2) * d(c) + d(x1) * d(s)).to(dt))
        print(dt, "neox" if neox else "gptj", "mismatches  A(per-op):", int((A != ref).sum()), " B(fp32):", int((B != ref).sum()),
              " D(fp64):", int((Dd != ref).sum()), "of", ref.numel())
