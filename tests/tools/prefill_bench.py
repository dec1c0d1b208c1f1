"""Prefill / mixed-batch kernel on the box: ours (tcgen05, FP8 KV) beside vLLM's Triton unified_attention and FlashInfer's
TRT-LLM-gen context kernel on the same paged FP8 cache, same inputs.  CUDA-graph replay, CUDA events, us per launch.

    python tests/tools/prefill_bench.py [--json out.json]

Useful FLOPs = 4 * D * Hq * sum over query tokens of their causal length (the two GEMMs, masked part not counted).
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
if os.path.isdir(os.path.join(os.path.dirname(__file__), "..", "..", "baseline", "_fi")):
    os.environ.setdefault("FLASHINFER_WORKSPACE_BASE", os.path.abspath(os.path.join(os.path.dirname(__file__), "..", "..", "baseline", "_fi")))

import ht_vllm_omni_b200.ops as ops  # noqa: E402
from tests.tools.sm100_reference import _graph_time  # noqa: E402

HQ, HKV, D, BS = 32, 4, 128, 16
WORK = {
    "prefill_8k": dict(ctx=[0], q=[8192]),
    "chunk_2k_on_30k": dict(ctx=[30720], q=[2048]),
    "8x512_on_4k": dict(ctx=[4096] * 8, q=[512] * 8),
    "verify_64x4_on_8k": dict(ctx=[8192] * 64, q=[4] * 64),
    "mixed_48dec_4x1k": dict(ctx=[8192] * 48 + [2048] * 4, q=[1] * 48 + [1024] * 4),
}


def problem(w, dev):
    ctx, ql = np.asarray(w["ctx"]), np.asarray(w["q"])
    lens = ctx + ql
    R, T = len(lens), int(ql.sum())
    nblk = [-(-int(l) // BS) for l in lens]
    nb = sum(nblk) + 4
    perm = torch.randperm(nb, generator=torch.Generator().manual_seed(5))[: sum(nblk)].to(torch.int32)
    bt = torch.zeros(R, max(nblk), dtype=torch.int32)
    o = 0
    for r in range(R):
        bt[r, : nblk[r]] = perm[o:o + nblk[r]]
        o += nblk[r]
    g = torch.Generator(device=dev).manual_seed(11)
    phys = torch.randint(0, 0x78, (nb, 2, HKV, BS, D), dtype=torch.uint8, device=dev, generator=g)
    phys |= torch.randint(0, 2, (nb, 2, HKV, BS, D), dtype=torch.uint8, device=dev, generator=g) << 7
    q = torch.randn(T, HQ, D, device=dev, generator=g).to(torch.bfloat16)
    qsl = torch.from_numpy(np.concatenate([[0], np.cumsum(ql)]).astype(np.int32)).to(dev)
    flops = 4.0 * D * HQ * float(sum(int(c) * int(n) + n * (n + 1) / 2 for c, n in zip(ctx, ql)))
    return dict(ctx=ctx, ql=ql, lens=torch.from_numpy(lens.astype(np.int32)).to(dev), bt=bt.to(dev), phys=phys, q=q, qsl=qsl, R=R, T=T,
                flops=flops, max_q=int(ql.max()), max_kv=int(lens.max()))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--json", default=None)
    ap.add_argument("--only", default=None)
    ap.add_argument("--ours-only", action="store_true", help="skip the comparison arms (ncu captures)")
    a = ap.parse_args()
    dev = torch.device("cuda")
    ks = torch.tensor([4.5 / 448.0], device=dev); vs = torch.tensor([4.5 / 448.0], device=dev)
    sm = D ** -0.5
    res = {}
    for name, w in WORK.items():
        if a.only and a.only != name:
            continue
        P = problem(w, dev)
        kc, vc = P["phys"].permute(0, 1, 3, 2, 4).unbind(1)
        row = {"requests": P["R"], "tokens": P["T"], "useful_gflop": P["flops"] / 1e9}
        # ---- ours: device-routed as the backend does (decode kernel for q_len 1, prefill kernel for the rest) ----
        out = torch.zeros_like(P["q"])
        tp = torch.zeros(P["R"] + 1, dtype=torch.int32, device=dev)
        has_dec = bool((P["ql"] < 2).any())
        torch.ops.b200attn.prefill_plan(P["qsl"], tp, HQ, HKV, 2)
        if has_dec:
            from ht_vllm_omni_b200.backend import expand_query_rows

            lens_t, rmap = expand_query_rows(P["qsl"], P["lens"], P["T"])
            short = (P["qsl"][1:] - P["qsl"][:-1]) < 2
            lens_t = torch.where(short[rmap.long()], lens_t, torch.zeros_like(lens_t)).contiguous()
            ws = ops.DecodeWorkspace(dev, P["T"], torch.uint8, head_dim=D, group_size=HQ // HKV)
            ws.make_plan(lens_t, HQ, HKV, P["bt"], BS, rmap)

        def ours():
            if has_dec:
                ws.decode(P["q"], kc, vc, P["bt"], lens_t, out, ks, vs, sm, row_map=rmap, fresh_tokens=1)
            torch.ops.b200attn.paged_prefill(P["q"], kc, vc, P["bt"], P["lens"], P["qsl"], tp, out, ks, vs, sm, 0)

        us, _ = _graph_time(ours, 1)
        row["ours"] = {"us": us, "tflops": P["flops"] / us / 1e6}
        if a.ours_only:
            res[name] = row
            print(name, json.dumps(row), flush=True)
            continue
        # ---- vLLM Triton unified_attention ----
        try:
            from vllm.v1.attention.ops.triton_unified_attention import unified_attention
            from vllm.v1.kv_cache_interface import KVQuantMode

            ref = torch.zeros_like(P["q"])
            k8, v8 = kc.view(torch.float8_e4m3fn), vc.view(torch.float8_e4m3fn)
            kd, vd = ks.reshape(1, 1).expand(P["R"], HKV), vs.reshape(1, 1).expand(P["R"], HKV)

            def tri():
                unified_attention(P["q"], k8, v8, ref, P["qsl"], P["max_q"], P["lens"], P["max_kv"], sm, True, (-1, -1), P["bt"], 0.0,
                                  None, kd, vd, kv_quant_mode=KVQuantMode.FP8_PER_TENSOR)

            us_t, _ = _graph_time(tri, 1)
            row["vllm_triton"] = {"us": us_t, "tflops": P["flops"] / us_t / 1e6, "max_abs_vs_ours": float((ref.float() - out.float()).abs().max())}
        except Exception as e:  # noqa: BLE001
            torch.cuda.synchronize()
            row["vllm_triton"] = {"error": f"{type(e).__name__}: {str(e)[:200]}"}
        # ---- FlashInfer TRT-LLM-gen context kernel (vLLM's default on sm100; fp8 query as vLLM feeds it with an fp8 cache) ----
        for qf8 in (True, False):
            key = "flashinfer_trtllm_q_fp8" if qf8 else "flashinfer_trtllm_q_bf16"
            try:
                from flashinfer.prefill import trtllm_batch_context_with_kv_cache

                wsb = torch.zeros(394 * 1024 * 1024, dtype=torch.uint8, device=dev)
                fo = torch.zeros_like(P["q"])
                q_scale = float(P["q"].float().abs().max()) / 448.0 if qf8 else 1.0
                qq = (P["q"].float() / q_scale).clamp(-448, 448).to(torch.float8_e4m3fn) if qf8 else P["q"]
                cache = P["phys"].view(torch.float8_e4m3fn)
                cum_kv = torch.cat([torch.zeros(1, dtype=torch.int32, device=dev), P["lens"].cumsum(0).to(torch.int32)])

                def fi():
                    trtllm_batch_context_with_kv_cache(query=qq, kv_cache=cache, workspace_buffer=wsb, block_tables=P["bt"], seq_lens=P["lens"],
                                                       max_q_len=P["max_q"], max_kv_len=P["max_kv"], bmm1_scale=sm * q_scale * float(ks),
                                                       bmm2_scale=float(vs), batch_size=P["R"], cum_seq_lens_q=P["qsl"], cum_seq_lens_kv=cum_kv,
                                                       out=fo, kv_layout="HND")

                us_f, _ = _graph_time(fi, 1)
                row[key] = {"us": us_f, "tflops": P["flops"] / us_f / 1e6, "max_abs_vs_ours": float((fo.float() - out.float()).abs().max())}
            except Exception as e:  # noqa: BLE001
                torch.cuda.synchronize()
                row[key] = {"error": f"{type(e).__name__}: {str(e)[:200]}"}
        res[name] = row
        print(name, json.dumps(row), flush=True)
        del P
        torch.cuda.empty_cache()
    if a.json:
        with open(a.json, "w") as f:
            json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
