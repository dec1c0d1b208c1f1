"""Where the prefill kernel's CTAs wait: per-role cycle counters of the debug instantiation (b200attn_debug_set_timeline_buffer).
Slots per CTA: softmax warp 0: [0] wait S ready [1] wait P buffer free [2] half-row exchange barrier; producer: [4] wait K slot
free [5] wait V slot free; MMA: [8] wait K tile [9] wait S buffer free [10] wait V tile [11] wait P ready; [12] CTA cycles,
[13] has work, [14] kv tiles.   python tests/tools/prefill_profile.py [workload]"""
import ctypes
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import ht_vllm_omni_b200.ops as ops  # noqa: E402
from tests.tools.prefill_bench import BS, D, HKV, HQ, WORK, problem  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "prefill_8k"
dev = torch.device("cuda")
P = problem(WORK[name], dev)
kc, vc = P["phys"].permute(0, 1, 3, 2, 4).unbind(1)
ks = torch.tensor([4.5 / 448.0], device=dev); vs = torch.tensor([4.5 / 448.0], device=dev)
out = torch.zeros_like(P["q"])
tp = torch.zeros(P["R"] + 1, dtype=torch.int32, device=dev)
torch.ops.b200attn.prefill_plan(P["qsl"], tp, HQ, HKV, 1)
tpt = ops.prefill_tile_tokens(HQ, HKV)
grid = (-(-P["T"] // tpt) + P["R"]) * HKV
buf = torch.zeros(grid * 16, dtype=torch.int64, device=dev)
lib = ops.load_library()
for _ in range(2):
    torch.ops.b200attn.paged_prefill(P["q"], kc, vc, P["bt"], P["lens"], P["qsl"], tp, out, ks, vs, D ** -0.5, 0)
torch.cuda.synchronize()
variant = int(sys.argv[2]) if len(sys.argv) > 2 else 0
lib.b200attn_set_option(b"debug", variant)
lib.b200attn_debug_set_timeline_buffer(ctypes.c_void_p(buf.data_ptr()))
torch.ops.b200attn.paged_prefill(P["q"], kc, vc, P["bt"], P["lens"], P["qsl"], tp, out, ks, vs, D ** -0.5, 0)
torch.cuda.synchronize()
lib.b200attn_debug_set_timeline_buffer(None)
lib.b200attn_set_option(b"debug", 0)
d = buf.cpu().numpy().reshape(grid, 16).astype(np.float64)
w = d[d[:, 13] > 0]
w = w[w[:, 14] >= 8]
n = w[:, 14]
print(f"{name} variant {variant}: {len(w)} CTAs with >= 8 kv tiles; cycles per kv tile (mean over CTAs)")
print(f"  CTA total            {np.mean(w[:, 12] / n):8.0f}")
for lbl, k in (("softmax: wait S", 0), ("softmax: wait P free", 1), ("softmax: exchange bar", 2), ("producer: wait K slot", 4),
               ("producer: wait V slot", 5), ("mma: wait K tile", 8), ("mma: wait S free", 9), ("mma: wait V tile", 10), ("mma: wait P ready", 11)):
    print(f"  {lbl:22s} {np.mean(w[:, k] / n):8.0f}")
# per-CTA fixed cost vs per-tile cost: least-squares line through (kv tiles, CTA cycles) over every CTA that had work
allw = d[d[:, 13] > 0]
A = np.stack([np.ones(len(allw)), allw[:, 14]], axis=1)
(a_fix, b_tile), *_ = np.linalg.lstsq(A, allw[:, 12], rcond=None)
print(f"  fit over {len(allw)} CTAs: cycles = {a_fix:.0f} + {b_tile:.0f} x kv tiles   (sum of CTA cycles / 148 SMs = {allw[:, 12].sum() / 148:.0f})")
for lo, hi in ((1, 2), (4, 5), (16, 17), (32, 33), (63, 65)):
    sel = allw[(allw[:, 14] >= lo) & (allw[:, 14] < hi)]
    if len(sel):
        print(f"    CTAs with {lo}-{hi - 1} kv tiles: {len(sel):4d}, mean cycles {sel[:, 12].mean():8.0f}")
