"""First-light diagnostics on a B200: runs a ladder of small cases through the C ABI and prints error magnitudes
instead of bare pass/fail, so one gpurun round trip says which piece (write / decode / merge / plan) is off."""
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import numpy as np
import torch

import ht_vllm_omni_b200.ops as ops  # noqa: F401  (registers torch.ops.b200attn)
from ht_vllm_omni_b200.ops import DecodeWorkspace
from tests.helpers import Case

torch.cuda.init()
print("device", torch.cuda.get_device_name(0), "sms", ops.load_library().b200attn_sm_count(), flush=True)


def run(name, lens, Hq, Hkv, kv, act="bf16", ctas=None, per_head=False, layout="HND", bs=16, check_write=True):
    c = Case(lens, Hq, Hkv, kv=kv, act=act, per_head=per_head, layout=layout, bs=bs).cuda()
    kc_o, vc_o = c.oracle_cache()
    msg = [f"{name:34s}"]
    try:
        torch.ops.b200attn.kv_write(c.d_key, c.d_value, c.kc, c.vc, c.d_slots, c.d_ks, c.d_vs)
        torch.cuda.synchronize()
        if check_write:
            km = int((c.cache_to_np(c.kc) != kc_o).sum()); vm = int((c.cache_to_np(c.vc) != vc_o).sum())
            msg.append(f"write mismatches K={km} V={vm}")
        c.load_cache_from_np(kc_o, vc_o)
        ws = DecodeWorkspace("cuda", c.B, torch.uint8 if c.fp8 else torch.bfloat16, num_ctas=ctas)
        out = torch.full((c.B, Hq, c.D), float("nan"), dtype=c.d_q.dtype, device="cuda")
        lse = torch.zeros(c.B, Hq, dtype=torch.float32, device="cuda")
        ws.make_plan(c.d_lens, Hq, Hkv, c.d_bt, c.bs)
        ws.decode(c.d_q, c.kc, c.vc, c.d_bt, c.d_lens, out, c.d_ks, c.d_vs, c.sm_scale, lse)
        torch.cuda.synchronize()
        got = out.float().cpu().numpy()
        ref = c.oracle_decode(kc_o, vc_o)
        ex = c.oracle_decode(kc_o, vc_o, "exact")
        valid = c.lens > 0
        err_ref = np.abs(got[valid] - ref[valid]); err_ex = np.abs(got[valid] - ex[valid])
        msg.append(f"ctas={ws.num_ctas} |cuda-oracle|max={np.nanmax(err_ref):.2e} |cuda-exact|max={np.nanmax(err_ex):.2e} "
                   f"|oracle-exact|max={np.abs(ref[valid]-ex[valid]).max():.2e} nan={int(np.isnan(got[valid]).sum())}")
        if np.isnan(got[valid]).any() or np.nanmax(err_ref) > 2e-2:
            per_seq = np.nanmax(np.where(np.isnan(got), 9.0, np.abs(got - ref)), axis=(1, 2))
            msg.append("per-seq err " + np.array2string(per_seq, precision=3))
            plan = ws.plan[:8].cpu().numpy(); msg.append(f"plan hdr {plan}")
    except Exception as e:  # noqa: BLE001
        msg.append(f"EXC {type(e).__name__}: {e}")
    print(" | ".join(msg), flush=True)


t0 = time.time()
run("fp8 1seq L=16 G8", [16], 8, 1, "fp8", ctas=1)
run("fp8 1seq L=64 G8", [64], 8, 1, "fp8", ctas=1)
run("fp8 1seq L=37 G8 (partial)", [37], 8, 1, "fp8", ctas=1)
run("fp8 1seq L=1000 1cta", [1000], 8, 1, "fp8", ctas=1)
run("fp8 1seq L=1000 4cta (merge)", [1000], 8, 1, "fp8", ctas=4)
run("fp8 ragged 5seq Hkv2 7cta", [1, 37, 100, 64, 513], 16, 2, "fp8", ctas=7)
run("fp8 ragged default grid", [1, 37, 100, 64, 513, 0, 2048], 16, 2, "fp8")
run("fp8 per-head scale", [300, 700], 16, 2, "fp8", per_head=True, ctas=5)
run("fp8 G=16", [300, 700], 16, 1, "fp8", ctas=3)
run("fp8 G=4", [300, 700], 8, 2, "fp8", ctas=3)
run("fp8 G=32 (2 passes)", [300, 200], 32, 1, "fp8", ctas=3)
run("fp8 fp16 act", [300, 700], 16, 2, "fp8", act="fp16", ctas=5)
run("fp8 NHD layout", [300, 700], 16, 2, "fp8", layout="NHD", ctas=5)
run("fp8 block_size 32", [300, 700], 16, 2, "fp8", bs=32, ctas=5)
run("bf16 1seq L=64", [64], 8, 1, "bf16", ctas=1)
run("bf16 1seq L=37", [37], 8, 1, "bf16", ctas=1)
run("bf16 ragged 7cta", [1, 37, 100, 64, 513], 16, 2, "bf16", ctas=7)
run("bf16 default grid", [1, 37, 100, 64, 513, 0, 2048], 16, 2, "bf16")
run("fp16 kv+act", [300, 700], 16, 2, "fp16", act="fp16", ctas=5)
run("bf16 G=16", [300, 700], 16, 1, "bf16", ctas=3)
run("fp8 B=64 L=4096 (16,2)", [4096] * 64, 16, 2, "fp8", check_write=False)
print("total s", round(time.time() - t0, 1))
