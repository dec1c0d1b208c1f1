"""BASELINE configs[3], the north-star multi-GPU config: Qwen3-Omni thinker decode (Hq 32 / Hkv 4, B 32, 32k context, FP8 KV)
STRONG-scaled over the N GPUs of one box, as vLLM's TP layout shards it ($VLLM/model_executor/models/qwen3.py:78-95,
vllm_omni/model_executor/stage_configs/qwen3_omni_moe.yaml:27 `tensor_parallel_size`):

  N = 2, 4   KV heads split across ranks (Hkv/N KV heads + their query heads per rank), no data-path collective;
  N = 8      tp > Hkv: two ranks share a KV head.  vLLM would replicate the head (no gain over N = 4); we split its CONTEXT
             between the two ranks (every other 16-token block), all-gather the group's query heads (NCCL, 32 KiB),
             decode the half context for all 8 heads returning (o, lse), and combine through NVLink peer memory
             (parallel.PeerExchange: symmetric-memory barrier + peer-load LSE merge) -- all inside one CUDA graph.

Called by bench.py (--gpus N > 1) on every rank; returns the `sharded` block (rank 0's copy is printed).  Times are CUDA
events around graph replays, max over ranks; efficiency = t_1gpu / (N * t_N).
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist

import ht_vllm_omni_b200.ops as ops
import ht_vllm_omni_b200.parallel as PAR

B, L, HQ, HKV, D, BS, LAYERS = 32, 32768, 32, 4, 128, 16, 2


def _time(body, dev, iters=10):
    """us per layer, max over ranks; `body` launches LAYERS layers and is captured into one CUDA graph."""
    side = torch.cuda.Stream(device=dev)
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        body(); body()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize(); dist.barrier()
    run, mode = body, "eager"
    try:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            body()
        run, mode = g.replay, "graph"
    except Exception:  # noqa: BLE001
        torch.cuda.synchronize()
    for _ in range(3):
        run()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        run()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) * 1e3 / iters / LAYERS], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()), mode


def run_sharded(dev, rank, world, lib):
    if world not in (2, 4, 8):
        return {"error": f"strong-scaled thinker block is defined for 2, 4 or 8 GPUs, not {world}"}
    gen = torch.Generator(device=dev).manual_seed(4242)     # SAME inputs on every rank: shards are slices of one problem
    nblk = L // BS
    nb = B * nblk + 8
    bt = torch.randperm(nb, generator=torch.Generator().manual_seed(7))[: B * nblk].to(torch.int32).reshape(B, nblk).to(dev)
    lens = torch.full((B,), L, dtype=torch.int32, device=dev)
    ks = torch.full((HKV,), 4.5 / 448.0, device=dev); vs = torch.full((HKV,), 4.5 / 448.0, device=dev)
    q = torch.randn(B, HQ, D, device=dev, generator=gen).to(torch.bfloat16)
    full = []
    for _ in range(LAYERS):
        ph = torch.randint(0, 0x78, (nb, 2, HKV, BS, D), dtype=torch.uint8, device=dev, generator=gen)
        ph |= (torch.randint(0, 2, (nb, 2, HKV, BS, D), dtype=torch.uint8, device=dev, generator=gen) << 7)
        full.append(ph)
    dec_bytes_full = 2 * B * L * HKV * D + 2 * B * HQ * D * 2 + B * nblk * 4 + 4 * B

    # ---- 1 GPU: the whole problem (every rank runs it; max over ranks is the baseline) ----
    ws1 = ops.DecodeWorkspace(dev, B, torch.uint8, head_dim=D, group_size=HQ // HKV)
    ws1.make_plan(lens, HQ, HKV, bt, BS)
    ref = torch.zeros(LAYERS, B, HQ, D, dtype=torch.bfloat16, device=dev)
    planes_full = [tuple(ph.permute(0, 1, 3, 2, 4).unbind(1)) for ph in full]

    def one_gpu():
        for l in range(LAYERS):
            ws1.decode(q, planes_full[l][0], planes_full[l][1], bt, lens, ref[l], ks, vs, D ** -0.5)

    t1, mode1 = _time(one_gpu, dev)

    # ---- this rank's shard: its KV head(s) in a compact cache of its own (what a TP rank allocates) ----
    kvh_per_rank = max(1, HKV // world)
    replicas = max(1, world // HKV)                       # ranks sharing one KV head (2 at N = 8)
    kv0 = (rank // replicas) * kvh_per_rank
    hq_per_kv = HQ // HKV
    shard = [ph[:, :, kv0:kv0 + kvh_per_rank].contiguous() for ph in full]
    del full, planes_full
    torch.cuda.empty_cache()
    planes = [tuple(ph.permute(0, 1, 3, 2, 4).unbind(1)) for ph in shard]
    ks_l, vs_l = ks[kv0:kv0 + kvh_per_rank].contiguous(), vs[kv0:kv0 + kvh_per_rank].contiguous()
    hq0 = kv0 * hq_per_kv
    hg = kvh_per_rank * hq_per_kv                          # query heads attending to this rank's KV head(s)
    q_group = q[:, hq0:hq0 + hg].contiguous()
    res = {"config": {"model": "Qwen3-Omni thinker attention shapes", "batch": B, "context": L, "num_q_heads": HQ, "num_kv_heads": HKV,
                      "head_dim": D, "block_size": BS, "kv_cache_dtype": "fp8", "layers_resident": LAYERS},
           "scaling": "strong", "n_gpus": world, "t_1gpu_us_per_layer": t1, "gbs_1gpu": dec_bytes_full / t1 / 1e3}

    if replicas == 1:
        # ---- head sharding: no exchange ----
        ws = ops.DecodeWorkspace(dev, B, torch.uint8, head_dim=D, group_size=hq_per_kv)
        ws.make_plan(lens, hg, kvh_per_rank, bt, BS)
        out = torch.zeros(LAYERS, B, hg, D, dtype=torch.bfloat16, device=dev)

        def layers():
            for l in range(LAYERS):
                ws.decode(q_group, planes[l][0], planes[l][1], bt, lens, out[l], ks_l, vs_l, D ** -0.5)

        tn, mode = _time(layers, dev)
        gathered = torch.empty(world * B, hg, D, dtype=torch.bfloat16, device=dev)

        def layers_ag():   # + the optional NCCL all-gather of head-sharded outputs (north star: "only when the layer runs head-sharded")
            for l in range(LAYERS):
                ws.decode(q_group, planes[l][0], planes[l][1], bt, lens, out[l], ks_l, vs_l, D ** -0.5)
                dist.all_gather_into_tensor(gathered, out[l])

        tn_ag, _ = _time(layers_ag, dev)
        err = float((out.float() - ref[:, :, hq0:hq0 + hg].float()).abs().max())
        res.update({"layout": f"tp{world}: {kvh_per_rank} KV head(s) + {hg} query heads per rank, no data-path collective",
                    "t_us_per_layer": tn, "exchange_us_per_layer": 0.0, "with_nccl_allgather_us_per_layer": tn_ag,
                    "allgather_bytes_per_layer": world * B * hg * D * 2})
    else:
        # ---- tp > Hkv: sequence split inside the replica group + peer-memory combine ----
        groups = [dist.new_group(list(range(g * replicas, (g + 1) * replicas))) for g in range(world // replicas)]
        grp = groups[rank // replicas]
        rr = rank % replicas
        bt_l = bt[:, rr::replicas].contiguous()                                   # every `replicas`-th 16-token block
        lens_l = PAR.dcp_local_seq_lens(lens, replicas, rr, BS)
        hq_local = hg // replicas
        q_local = q_group[:, rr * hq_local:(rr + 1) * hq_local].contiguous()
        ws = ops.DecodeWorkspace(dev, B, torch.uint8, head_dim=D, group_size=hq_per_kv)
        ws.make_plan(lens_l, hg, 1, bt_l, BS)
        px = PAR.PushExchange(B, hq_local, D, torch.bfloat16, dev, group=grp)
        outs = [None] * LAYERS

        def layers(exchange=True):
            for l in range(LAYERS):
                qg = px.gather_q(q_local) if exchange else q_group
                o_buf, lse_buf = px.partial_buffers(B)
                ws.decode(qg, planes[l][0], planes[l][1], bt_l, lens_l, o_buf, ks_l, vs_l, D ** -0.5, lse=lse_buf)
                if exchange:
                    outs[l] = px.combine(B)

        tn, mode = _time(layers, dev)
        td, _ = _time(lambda: layers(False), dev)
        layers()
        torch.cuda.synchronize()
        got = torch.stack(outs)
        h0 = hq0 + rr * hq_local
        err = float((got.float() - ref[:, :, h0:h0 + hq_local].float()).abs().max())
        # the same layers with the round-1 exchange for comparison: NCCL all-gather of the query heads + symmetric-memory
        # barrier + peer-LOAD merge (parallel.PeerExchange)
        pl = PAR.PeerExchange(B, hg, D, torch.bfloat16, dev, group=grp)
        qg_buf = torch.empty(replicas * B, hq_local, D, dtype=torch.bfloat16, device=dev)

        def layers_pull():
            for l in range(LAYERS):
                dist.all_gather_into_tensor(qg_buf, q_local, group=grp)
                qg = qg_buf.view(replicas, B, hq_local, D).permute(1, 0, 2, 3).reshape(B, hg, D)
                o_buf, lse_buf = pl.buffers()
                ws.decode(qg, planes[l][0], planes[l][1], bt_l, lens_l, o_buf, ks_l, vs_l, D ** -0.5, lse=lse_buf)
                pl.combine(B)

        try:
            t_pull, _ = _time(layers_pull, dev)
        except Exception:  # noqa: BLE001
            t_pull = None
        res.update({"layout": f"tp{world}: KV head x {replicas}-way sequence split (16-token interleave); per rank {hq_local} query heads, "
                              f"half the context of one KV head; query heads and (o, lse) partial slices PUSHED over NVLink peer memory "
                              f"(two flag-synchronised launches per layer, no collective)",
                    "t_us_per_layer": tn, "decode_only_us_per_layer": td, "exchange_us_per_layer": tn - td,
                    "nccl_allgather_plus_peer_load_merge_us_per_layer": t_pull})
    e = torch.tensor([err], device=dev, dtype=torch.float64)
    dist.all_reduce(e, op=dist.ReduceOp.MAX)
    res.update({"efficiency": t1 / (world * tn), "speedup": t1 / tn, "gbs_aggregate": dec_bytes_full / tn / 1e3,
                "max_abs_vs_unsharded": float(e.item()), "tolerance": 1e-2, "timing": f"{mode} replay, CUDA events, max over ranks"})
    return res
