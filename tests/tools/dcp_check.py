"""Sequence-split ("DCP") exchange on real GPUs over NCCL (tp > Hkv): the ranks of a replica group share ONE
KV head, each keeps every dcp-th run of 16 tokens, runs the decode kernel over its slice for ALL the group's query heads
(returning o and lse), and the partials are combined with all-gather + torch.ops.b200attn.merge_lse.  Checked against the
same kernel run unsharded on the full context (every rank builds it), max-abs <= 1e-2 (fp8 KV tolerance).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tests/tools/dcp_check.py
"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import numpy as np
import torch
import torch.distributed as dist
import ht_vllm_omni_b200.ops as ops
import ht_vllm_omni_b200.parallel as P

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
B, Hl, D, bs, IL = 16, 8 // world, 128, 16, 16
Hg = Hl * world
g = torch.Generator().manual_seed(7)                       # same stream on every rank
lens = torch.randint(1, 4096, (B,), generator=g).to(torch.int32)
lens[0], lens[1] = 5, 4096
Lmax = int(lens.max())
K = torch.randn(B, Lmax, 1, D, generator=g).bfloat16()
V = torch.randn(B, Lmax, 1, D, generator=g).bfloat16()
q_all = torch.randn(B, Hg, D, generator=g).bfloat16()
ks = torch.full((1,), 4.5 / 448, device=dev); vs = torch.full((1,), 4.5 / 448, device=dev)


def build(positions_per_seq):
    """paged fp8 cache holding, for sequence s, the tokens positions_per_seq[s] in that order; returns planes, bt, lens"""
    n = [len(p) for p in positions_per_seq]
    nblk = max(1, -(-max(n) // bs))
    nb = B * nblk + 1
    bt = torch.randperm(nb - 1, generator=torch.Generator().manual_seed(3))[: B * nblk].to(torch.int32).reshape(B, nblk)
    phys = torch.zeros((nb, 2, 1, bs, D), dtype=torch.uint8, device=dev)
    kc, vc = phys.permute(0, 1, 3, 2, 4).unbind(1)
    for s in range(B):
        if n[s] == 0:
            continue
        idx = torch.tensor(positions_per_seq[s])
        j = torch.arange(n[s])
        slots = (bt[s, j // bs].to(torch.int64) * bs + j % bs).to(dev)
        torch.ops.b200attn.kv_write(K[s, idx].to(dev), V[s, idx].to(dev), kc, vc, slots, ks, vs)
    return kc, vc, bt.to(dev), torch.tensor(n, dtype=torch.int32, device=dev)


def decode(kc, vc, bt, ln, q, want_lse):
    ws = ops.DecodeWorkspace(dev, B, torch.uint8, group_size=q.shape[1])
    out = torch.zeros(B, q.shape[1], D, dtype=torch.bfloat16, device=dev)
    lse = torch.full((B, q.shape[1]), float("-inf"), dtype=torch.float32, device=dev) if want_lse else None
    ws.make_plan(ln, q.shape[1], 1, bt, kc.shape[1])
    ws.decode(q, kc, vc, bt, ln, out, ks, vs, D ** -0.5, lse=lse)
    return out, lse


# unsharded truth on every rank
full_out, _ = decode(*build([list(range(int(l))) for l in lens]), q_all.to(dev), False)
# sequence-split: this rank's slice
mine = [P.dcp_local_positions(int(l), world, rank, IL) for l in lens]
kc, vc, bt, ln = build(mine)
assert torch.equal(ln.cpu(), P.dcp_local_seq_lens(lens, world, rank, IL))
q_local = q_all[:, rank * Hl:(rank + 1) * Hl].contiguous().to(dev)
q_group = P.dcp_gather_queries(q_local)                    # NCCL all-gather of the group's query heads
assert torch.equal(q_group.cpu(), q_all)
o_part, lse_part = decode(kc, vc, bt, ln, q_group, True)
evs = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
got = P.dcp_combine(o_part, lse_part)                      # NCCL all-gathers + merge_lse kernel
torch.cuda.synchronize(); dist.barrier()
evs[0].record()
for _ in range(20):
    got = P.dcp_combine(o_part, lse_part)
evs[1].record(); torch.cuda.synchronize()
err = (got.float() - full_out[:, rank * Hl:(rank + 1) * Hl].float()).abs().max().item()
t = torch.tensor([err], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    print(f"dcp{world} over NCCL: max |sequence-split - unsharded| = {t.item():.5f} (tolerance 1e-2); "
          f"combine (2 all-gathers + merge_lse) {evs[0].elapsed_time(evs[1]) / 20 * 1e3:.1f} us per layer", flush=True)
assert t.item() <= 1e-2, t.item()

# ---- the same exchange over NVLink peer memory: decode writes its partial into a symmetric buffer, barrier, peer-load merge ----
px = P.PeerExchange(B, Hg, D, torch.bfloat16, dev)
ws = ops.DecodeWorkspace(dev, B, torch.uint8, group_size=Hg)
ws.make_plan(ln, Hg, 1, bt, kc.shape[1])


def layer():
    o_buf, lse_buf = px.buffers()           # NOT pre-filled: rows without tokens on this rank (sequence 0 has 5 tokens, all on
    ws.decode(q_group, kc, vc, bt, ln, o_buf, ks, vs, D ** -0.5, lse=lse_buf)   # rank 0) come back as (0, -inf) from the kernel
    return px.combine(B)


got2 = layer()
for _ in range(3):
    got2 = layer()
torch.cuda.synchronize(); dist.barrier()
e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
e0.record()
for _ in range(20):
    got2 = layer()
e1.record()
for _ in range(20):
    o_buf, lse_buf = px.buffers()
    ws.decode(q_group, kc, vc, bt, ln, o_buf, ks, vs, D ** -0.5, lse=lse_buf)
e2.record(); torch.cuda.synchronize()
err2 = (got2.float() - full_out[:, rank * Hl:(rank + 1) * Hl].float()).abs().max().item()
t2 = torch.tensor([err2], device=dev); dist.all_reduce(t2, op=dist.ReduceOp.MAX)
if rank == 0:
    comb = (e0.elapsed_time(e1) - e1.elapsed_time(e2)) / 20 * 1e3
    print(f"dcp{world} over NVLink peer memory: max |sequence-split - unsharded| = {t2.item():.5f}; decode+combine "
          f"{e0.elapsed_time(e1) / 20 * 1e3:.1f} us per layer, of which barrier + peer-load merge {comb:.1f} us", flush=True)
assert t2.item() <= 1e-2, t2.item()
# ---- push exchange: query heads pushed to the peers' gather buffers, partial slices pushed to the peers' receive buffers ----
pp = P.PushExchange(B, Hl, D, torch.bfloat16, dev)


def layer_push():
    qg = pp.gather_q(q_local)
    o_p, l_p = pp.partial_buffers(B)
    ws.decode(qg, kc, vc, bt, ln, o_p, ks, vs, D ** -0.5, lse=l_p)
    return pp.combine(B)


got3 = layer_push()
assert torch.equal(pp.gathered[pp.slot, :B].cpu(), q_all), "pushed query heads differ from the all-gathered ones"
for _ in range(3):
    got3 = layer_push()
torch.cuda.synchronize(); dist.barrier()
# replayed as a CUDA graph too (flags are launch counters: nothing to reset between replays)
side = torch.cuda.Stream(device=dev); side.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(side):
    layer_push(); layer_push()
torch.cuda.current_stream().wait_stream(side); torch.cuda.synchronize(); dist.barrier()
gr = torch.cuda.CUDAGraph()
with torch.cuda.graph(gr):
    got3 = layer_push()
    got3b = layer_push()
for _ in range(3):
    gr.replay()
torch.cuda.synchronize(); dist.barrier()
e0, e1 = (torch.cuda.Event(enable_timing=True) for _ in range(2))
e0.record()
for _ in range(10):
    gr.replay()
e1.record(); torch.cuda.synchronize()
err3 = max((got3.float() - full_out[:, rank * Hl:(rank + 1) * Hl].float()).abs().max().item(),
           (got3b.float() - full_out[:, rank * Hl:(rank + 1) * Hl].float()).abs().max().item())
t3 = torch.tensor([err3], device=dev); dist.all_reduce(t3, op=dist.ReduceOp.MAX)
if rank == 0:
    print(f"dcp{world} over NVLink pushes (gather_q + decode + combine, graph replay): max |sequence-split - unsharded| = {t3.item():.5f}; "
          f"{e0.elapsed_time(e1) / 20 * 1e3:.1f} us per layer", flush=True)
assert t3.item() <= 1e-2, t3.item()
assert torch.equal(got3, got2), "push combine and pull combine disagree"
dist.barrier(); torch.cuda.synchronize()
sys.stdout.flush()
os._exit(0)
