"""Cost of the push exchange steps alone (no decode), per launch, CUDA-graph replay, max over ranks.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tests/tools/push_exchange_bench.py"""
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
lib = ops.load_library()
B, Hl, D = 32, 4, 128
px = P.PushExchange(B, Hl, D, torch.bfloat16, dev)
q_local = torch.randn(B, Hl, D, device=dev).to(torch.bfloat16)
N = 20


def bench(name, body):
    side = torch.cuda.Stream(device=dev); side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        body(); body()
    torch.cuda.current_stream().wait_stream(side); torch.cuda.synchronize(); dist.barrier()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(N):
            body()
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        g.replay()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) * 1e3 / 10 / N], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(f"{name}: {t.item():.2f} us per call", flush=True)


for pdl in (1, 0):
    lib.b200attn_set_option(b"pdl", pdl)
    if rank == 0:
        print(f"-- pdl={pdl}", flush=True)
    bench("gather_q", lambda: px.gather_q(q_local))
    bench("combine", lambda: px.combine(B))
    bench("gather_q + combine", lambda: (px.gather_q(q_local), px.combine(B)))
    out = torch.empty(B, Hl, D, dtype=torch.bfloat16, device=dev)
    bench("combine into a fixed out", lambda: px.combine(B, out))
    bench("empty-ish kernel (merge_lse of 1 part)", lambda: torch.ops.b200attn.merge_lse(px.o_part[None, :, :Hl].contiguous(), px.lse_part[None, :, :Hl].contiguous(), out, torch.empty(0, device=dev)))
dist.barrier(); torch.cuda.synchronize()
os._exit(0)
