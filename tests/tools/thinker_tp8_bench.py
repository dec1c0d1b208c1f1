"""BASELINE configs[3] at tp = 8 > Hkv = 4: one replica group (2 ranks sharing ONE KV head of the Qwen3-Omni thinker, 8 query
heads, B 32, 32k context, FP8 KV) with the sequence-split exchange over NVLink peer memory.  The four groups of an 8-GPU
box are independent, so two GPUs measure the per-layer time of the whole tp = 8 layout.  Also times the tp = 4 layout
(one rank, one KV head, whole context, no exchange) on rank 0 for the strong-scaling step 4 -> 8.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tests/tools/thinker_tp8_bench.py
"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import torch
import torch.distributed as dist
import ht_vllm_omni_b200.ops as ops
import ht_vllm_omni_b200.parallel as P

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
B, L, G, D, bs, LAYERS = 32, 32768, 8, 128, 16, 4   # even: the two exchange slots keep alternating across graph replays
ks = torch.full((1,), 0.01, device=dev); vs = torch.full((1,), 0.01, device=dev)
q = torch.randn(B, G, D, device=dev).bfloat16()


def caches(ctx):
    nblk = ctx // bs
    nb = B * nblk + 8
    bt = torch.randperm(nb, generator=torch.Generator().manual_seed(1))[: B * nblk].to(torch.int32).reshape(B, nblk).to(dev)
    cs = [torch.randint(0, 0x78, (nb, 2, 1, bs, D), dtype=torch.uint8, device=dev).permute(0, 1, 3, 2, 4).unbind(1) for _ in range(LAYERS)]
    return cs, bt, torch.full((B,), ctx, dtype=torch.int32, device=dev)


def time_us(body, iters=10):
    """us per layer, max over ranks; the layers are captured into one CUDA graph (falls back to eager launches)."""
    side = torch.cuda.Stream(); side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        body()
    torch.cuda.current_stream().wait_stream(side); torch.cuda.synchronize(); dist.barrier()
    run, mode = body, "eager"
    try:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            body()
        run, mode = g.replay, "graph"
    except Exception as e:  # noqa: BLE001
        if rank == 0:
            print(f"# graph capture failed ({type(e).__name__}: {str(e)[:120]}); eager launches", flush=True)
        torch.cuda.synchronize()
    for _ in range(3):
        run()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        run()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) * 1e3 / iters / LAYERS], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item(), mode


# ---- tp = 8: this rank holds every other run of 16 tokens: L / 2 tokens per sequence ----
cs, bt, ln = caches(L // world)
ws = ops.DecodeWorkspace(dev, B, torch.uint8, group_size=G)
ws.make_plan(ln, G, 1, bt, cs[0][0].shape[1])
px = P.PeerExchange(B, G, D, torch.bfloat16, dev)


def tp8_layers():
    for l in range(LAYERS):
        o_buf, lse_buf = px.buffers()
        ws.decode(q, cs[l][0], cs[l][1], bt, ln, o_buf, ks, vs, D ** -0.5, lse=lse_buf)
        px.combine(B)


def tp8_decode_only():
    for l in range(LAYERS):
        o_buf, lse_buf = px.buffers()
        ws.decode(q, cs[l][0], cs[l][1], bt, ln, o_buf, ks, vs, D ** -0.5, lse=lse_buf)


(t8, m8), (t8d, _) = time_us(tp8_layers), time_us(tp8_decode_only)
del cs
torch.cuda.empty_cache()
# ---- tp = 4: one rank, whole context, no exchange (every rank runs it; max over ranks) ----
cs4, bt4, ln4 = caches(L)
ws4 = ops.DecodeWorkspace(dev, B, torch.uint8, group_size=G)
ws4.make_plan(ln4, G, 1, bt4, cs4[0][0].shape[1])
out4 = torch.zeros(B, G, D, dtype=torch.bfloat16, device=dev)


def tp4_layers():
    for l in range(LAYERS):
        ws4.decode(q, cs4[l][0], cs4[l][1], bt4, ln4, out4, ks, vs, D ** -0.5)


t4, m4 = time_us(tp4_layers)
if rank == 0:
    kv_bytes = 2 * B * L * D   # one KV head, fp8
    print(f"thinker (B {B}, {L} ctx, fp8, one KV head / 8 q heads per replica group), us per layer (tp=4: {m4}, tp=8: {m8}):")
    print(f"  tp=4 (1 rank / KV head, no exchange): {t4:.1f} us  = {kv_bytes / t4 / 1e3:.0f} GB/s per KV head")
    print(f"  tp=8 (2 ranks / KV head, sequence split + peer-memory merge): {t8:.1f} us (decode alone {t8d:.1f}) = "
          f"{kv_bytes / t8 / 1e3:.0f} GB/s per KV head -> strong-scaling step 4->8: {t4 / t8:.2f}x of an ideal 2x", flush=True)
dist.barrier(); torch.cuda.synchronize()
sys.stdout.flush()
os._exit(0)
