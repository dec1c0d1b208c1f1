"""Ours vs the reference's own sm_100 kernels (FlashInfer TRT-LLM-gen, Triton unified_attention, FA4) on the same inputs.

    python tests/tools/ref_compare.py [--workloads c2,thinker,c2-bf16,b1-128k,ragged,tail] [--out gpurun_out/ref_compare]

One markdown row per (workload, arm): us per layer, algorithmic GB/s, ratio ours/arm, max-abs difference of the outputs.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import torch

import ht_vllm_omni_b200.ops as ops
from tests.tools import sm100_reference as R
from tests.tools.workloads import WORKLOADS, Problem


def time_ours(P, lib, pdl):
    w = P.w
    ws = ops.DecodeWorkspace(P.q.device, w["B"], P.kv_t, head_dim=w["D"], group_size=w["Hq"] // w["Hkv"])
    ws.make_plan(P.lens, w["Hq"], w["Hkv"], P.bt, w["bs"])
    torch.cuda.synchronize()
    lib.b200attn_set_option(b"pdl", pdl)

    def body():
        for l in range(P.NL):
            kc, vc = P.planes[l]
            ws.decode(P.q[l], kc, vc, P.bt, P.lens, P.out[l], P.k_scale, P.v_scale, P.sm_scale)

    us, graphed = R._graph_time(body, P.NL)
    lib.b200attn_set_option(b"pdl", 1)
    return {"us_per_layer": us, "gbs": P.dec_bytes / us / 1e3, "cuda_graph": graphed, "grid": ws.grid}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workloads", default="c2,thinker,c2-bf16,b1-128k,ragged,tail,c1")
    ap.add_argument("--out", default="gpurun_out/ref_compare")
    ap.add_argument("--arms", default="flashinfer,flashinfer_q8,triton,fa4,kv_write")
    a = ap.parse_args()
    dev = torch.device("cuda")
    lib = ops.load_library()
    results = {}
    lines = ["| workload | arm | us/layer | GB/s | ours(pdl off)/arm time | max abs vs ours |", "|---|---|---|---|---|---|"]
    for name in a.workloads.split(","):
        w = WORKLOADS[name]
        P = Problem(w, dev, fill="random")
        ours0 = time_ours(P, lib, 0)
        ours1 = time_ours(P, lib, 1)
        ours_out = P.out.clone()
        arms = R.compare_all(P.phys, P.q, P.bt, P.lens, P.sm_scale, P.k_scale, P.v_scale, P.dec_bytes, P.wr_bytes,
                             kn=P.kn, vn=P.vn, slots=P.step_slots, ours_out=ours_out, arms=tuple(a.arms.split(",")))
        # our kv_write for the write row
        def wr():
            for l in range(P.NL):
                torch.ops.b200attn.kv_write(P.kn[l], P.vn[l], P.planes[l][0], P.planes[l][1], P.step_slots, P.k_scale, P.v_scale)
        us_w, gw = R._graph_time(wr, P.NL)
        results[name] = {"ours_pdl_off": ours0, "ours_pdl_on": ours1, "ours_kv_write": {"us_per_layer": us_w, "gbs": P.wr_bytes / us_w / 1e3},
                         "dec_bytes": P.dec_bytes, **arms}
        lines.append(f"| {name} | ours decode (graph, PDL off) grid {ours0['grid']} | {ours0['us_per_layer']:.2f} | {ours0['gbs']:.0f} | 1.00 | - |")
        lines.append(f"| {name} | ours decode (graph, PDL on) | {ours1['us_per_layer']:.2f} | {ours1['gbs']:.0f} | {ours0['us_per_layer'] / ours1['us_per_layer']:.2f} | - |")
        lines.append(f"| {name} | ours kv_write | {us_w:.2f} | {P.wr_bytes / us_w / 1e3:.0f} | - | - |")
        for k, v in arms.items():
            if "error" in v:
                lines.append(f"| {name} | {k} | - | - | - | {v['error'][:160]} |")
            else:
                base = us_w if k.startswith("vllm_reshape") else ours0["us_per_layer"]
                ma = v.get("max_abs_vs_ours")
                lines.append(f"| {name} | {k} {v.get('q_dtype', '')}{v.get('split', '')} | {v['us_per_layer']:.2f} | {v['gbs']:.0f} | "
                             f"{base / v['us_per_layer']:.2f} | {'-' if ma is None else f'{ma:.2e}'} |")
        print("\n".join(lines[-(4 + len(arms)):]), flush=True)
        del P
        torch.cuda.empty_cache()
    os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
    with open(a.out + ".md", "w") as f:
        f.write("\n".join(lines) + "\n")
    with open(a.out + ".json", "w") as f:
        json.dump(results, f, indent=1)


if __name__ == "__main__":
    main()
