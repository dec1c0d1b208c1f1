"""Multi-GPU layout of the paged-decode path: KV-head sharding (vLLM's TP layout) and the sequence-split exchange
used when there are more ranks than KV heads.  One process per GPU, torch.distributed (NCCL on the
box, gloo in the CPU tests) for the plumbing; the arithmetic stays in libb200attn.

Head sharding needs NO collective inside attention: rank r owns KV heads [r*Hkv/tp, (r+1)*Hkv/tp) with their G query
heads, its own cache shard, and replicated block tables / seq_lens ($VLLM/model_executor/models/qwen3.py:78-95); the
row-parallel o_proj that follows consumes head-sharded outputs.  `all_gather_heads` is the optional NCCL all-gather the
north star names, for callers that want unsharded heads.

With tp > Hkv vLLM replicates each KV head on tp/Hkv ranks (qwen3.py:88-91), so KV bytes per GPU stop shrinking.  The
sequence-split ("DCP") remedy: the ranks of one replica group each keep every `dcp`-th token of the context
($VLLM/v1/worker/block_table.py:148-162), all-gather the group's query heads (tiny), run the same decode kernel over
their slice returning (o, lse), and combine by log-sum-exp (replaces cp_lse_ag_out_rs,
$VLLM/v1/attention/backends/flashinfer.py:1714-1738) with `torch.ops.b200attn.merge_lse`.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.distributed as dist


@dataclass(frozen=True)
class HeadShard:
    """What one TP rank owns.  Query heads are always split evenly; a KV head is replicated when tp > Hkv."""

    tp_size: int
    tp_rank: int
    total_q_heads: int
    total_kv_heads: int

    def __post_init__(self):
        if self.total_q_heads % self.total_kv_heads:
            raise ValueError("Hq must be a multiple of Hkv")
        if self.total_q_heads % self.tp_size:
            raise ValueError("Hq must be divisible by tp_size")
        if self.tp_size <= self.total_kv_heads:
            if self.total_kv_heads % self.tp_size:
                raise ValueError("Hkv must be divisible by tp_size")
        elif self.tp_size % self.total_kv_heads:
            raise ValueError("tp_size must be a multiple of Hkv when it exceeds it")
        if not 0 <= self.tp_rank < self.tp_size:
            raise ValueError("tp_rank out of range")

    @property
    def num_q_heads(self) -> int:
        return self.total_q_heads // self.tp_size

    @property
    def num_kv_heads(self) -> int:
        return max(1, self.total_kv_heads // self.tp_size)

    @property
    def replicas(self) -> int:
        """Ranks holding the same KV head (1 while tp <= Hkv): the size of a sequence-split group."""
        return max(1, self.tp_size // self.total_kv_heads)

    @property
    def replica_rank(self) -> int:
        return self.tp_rank % self.replicas

    @property
    def q_head_range(self) -> range:
        return range(self.tp_rank * self.num_q_heads, (self.tp_rank + 1) * self.num_q_heads)

    @property
    def kv_head_range(self) -> range:
        if self.tp_size <= self.total_kv_heads:
            return range(self.tp_rank * self.num_kv_heads, (self.tp_rank + 1) * self.num_kv_heads)
        h = self.tp_rank // self.replicas
        return range(h, h + 1)

    def replica_group_ranks(self) -> list[int]:
        base = (self.tp_rank // self.replicas) * self.replicas
        return list(range(base, base + self.replicas))


def all_gather_heads(out_local: torch.Tensor, group=None) -> torch.Tensor:
    """[T, Hq/tp, D] on every rank -> [T, Hq, D] (heads in rank order).  One all-gather, no other traffic."""
    world = dist.get_world_size(group)
    if world == 1:
        return out_local
    T, Hl, D = out_local.shape
    buf = torch.empty((world * T, Hl, D), dtype=out_local.dtype, device=out_local.device)   # rank-major concat along dim 0
    dist.all_gather_into_tensor(buf, out_local.contiguous(), group=group)
    return buf.view(world, T, Hl, D).permute(1, 0, 2, 3).reshape(T, world * Hl, D)


def dcp_local_seq_lens(seq_lens: torch.Tensor, dcp_size: int, dcp_rank: int, interleave: int = 1) -> torch.Tensor:
    """Tokens of a context are dealt round-robin in runs of `interleave` to the ranks of a sequence-split group:
    position p lives on rank (p // interleave) % dcp_size.  Returns how many of the first seq_lens[s] positions this rank
    holds (int32, same device, no host sync) — the `seq_lens` its local decode call gets."""
    L = seq_lens.to(torch.int64)
    round_len = dcp_size * interleave
    full = (L // round_len) * interleave
    rem = (L % round_len) - dcp_rank * interleave
    return (full + rem.clamp(0, interleave)).to(torch.int32)


def dcp_local_positions(seq_len: int, dcp_size: int, dcp_rank: int, interleave: int = 1) -> list[int]:
    """Host-side mirror of dcp_local_seq_lens: the global positions a rank holds, in local order."""
    return [p for p in range(seq_len) if (p // interleave) % dcp_size == dcp_rank]


def _cuda_merge_lse(parts_o, parts_lse, out, lse_out):
    torch.ops.b200attn.merge_lse(parts_o, parts_lse, out, lse_out)   # raises on CPU tensors: no CPU fallback


def dcp_gather_queries(q_local: torch.Tensor, group=None) -> torch.Tensor:
    """All-gather the replica group's query heads: [T, Hq_local, D] -> [T, dcp*Hq_local, D] (all share one KV head)."""
    return all_gather_heads(q_local, group)


def dcp_combine(out_part: torch.Tensor, lse_part: torch.Tensor, group=None, merge_fn=None,
                keep_local_heads: bool = True) -> torch.Tensor:
    """Combine the per-rank partial attention of a sequence-split group.

    out_part [T, Hg, D] / lse_part [T, Hg] fp32 (natural log; -inf where the rank held no token) are this rank's result
    for ALL Hg = dcp*Hq_local heads of the group over ITS slice of the context.  All-gathers both, merges by
    log-sum-exp (`merge_fn`, default the CUDA kernel), and returns this rank's own Hq_local heads (or all Hg).
    """
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    T, Hg, D = out_part.shape
    merge = merge_fn or _cuda_merge_lse
    parts_o = torch.empty((world, T, Hg, D), dtype=out_part.dtype, device=out_part.device)
    parts_lse = torch.empty((world, T, Hg), dtype=torch.float32, device=out_part.device)
    if world > 1:
        dist.all_gather_into_tensor(parts_o.view(world * T, Hg, D), out_part.contiguous(), group=group)
        dist.all_gather_into_tensor(parts_lse.view(world * T, Hg), lse_part.contiguous(), group=group)
    else:
        parts_o[0].copy_(out_part)
        parts_lse[0].copy_(lse_part)
    out = torch.empty((T, Hg, D), dtype=out_part.dtype, device=out_part.device)
    merge(parts_o, parts_lse, out, torch.empty(0, dtype=torch.float32, device=out_part.device))
    if not keep_local_heads:
        return out
    hl = Hg // world
    return out[:, rank * hl:(rank + 1) * hl]


class PeerExchange:
    """Sequence-split combine over NVLink peer memory instead of NCCL (SURVEY.md §8e remedy 1, "write partials peer-to-peer
    into a symmetric buffer and run Kernel 3"): every rank of the group decodes straight into its slot of a torch
    symmetric-memory buffer, one device-side barrier orders the writes, and `b200attn_merge_lse_peers` reads all ranks'
    partials with peer loads while it merges.  Two slots alternate, so a rank that runs ahead cannot overwrite a partial a
    peer is still reading (it would have to pass the NEXT barrier first, which that peer only joins after its merge).
    One launch + one barrier per layer replace two all-gathers, three allocations and a merge launch.
    """

    def __init__(self, max_rows: int, group_heads: int, head_dim: int, dtype: torch.dtype, device, group=None):
        import torch.distributed._symmetric_memory as symm_mem

        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        self.T, self.Hg, self.D, self.dtype = max_rows, group_heads, head_dim, dtype
        self.o = symm_mem.empty((2, max_rows, group_heads, head_dim), dtype=dtype, device=device)
        self.lse = symm_mem.empty((2, max_rows, group_heads), dtype=torch.float32, device=device)
        self.h_o = symm_mem.rendezvous(self.o, self.group)
        self.h_lse = symm_mem.rendezvous(self.lse, self.group)
        self.slot = 0

    def buffers(self) -> tuple[torch.Tensor, torch.Tensor]:
        """(out, lse) views of the NEXT slot: pass them to the local decode call so it writes the partial in place."""
        self.slot ^= 1
        return self.o[self.slot], self.lse[self.slot]

    def combine(self, num_rows: int, keep_local_heads: bool = True) -> torch.Tensor:
        """Barrier + peer-memory merge of the slot handed out by the last `buffers()` call."""
        from ._lib import DT_BF16, DT_FP16, check, load_library

        dev = self.o.device
        self.h_o.barrier(channel=self.slot)
        out = torch.empty((num_rows, self.Hg, self.D), dtype=self.dtype, device=dev)
        stream = torch.cuda.current_stream(dev)
        check(load_library().b200attn_merge_lse_peers(
            self.h_o.buffer_ptrs_dev, self.h_lse.buffer_ptrs_dev, self.world,
            self.slot * self.T * self.Hg * self.D, self.slot * self.T * self.Hg, out.data_ptr(), None, num_rows, self.Hg, self.D,
            DT_BF16 if self.dtype == torch.bfloat16 else DT_FP16, stream.cuda_stream), "b200attn_merge_lse_peers")
        if not keep_local_heads:
            return out
        hl = self.Hg // self.world
        return out[:, self.rank * hl:(self.rank + 1) * hl]


class PushExchange:
    """The sequence-split exchange of a replica group as two PUSH launches per layer around the decode, no collective, no
    barrier and no memory fence (csrc/peer_exchange.cu, "LL" words: every 8-byte store carries its own flag): `gather_q`
    replaces the all-gather of the group's query heads, `combine` the all-gather / reduce-scatter of partials + merge
    (cp_lse_ag_out_rs, $VLLM/v1/attention/backends/flashinfer.py:1714-1738).

        qg = px.gather_q(q_local)                       # [n, R * Hl, D]: every rank's heads, pushed by their owners
        o_part, lse_part = px.partial_buffers(n)        # hand these to the local decode (lse requested)
        ws.decode(qg, ..., o_part, ..., lse=lse_part)
        out = px.combine(n)                             # [n, Hl, D]: this rank's heads over the whole context

    Every rank of the group must make the same calls with the same n.  Receive buffers are torch symmetric-memory
    allocations; flags are launch counters, so the object is CUDA-graph safe (capture and replay see the same addresses,
    nothing is reset)."""

    def __init__(self, max_rows: int, heads_local: int, head_dim: int, dtype: torch.dtype, device, group=None):
        import torch.distributed._symmetric_memory as symm_mem

        from ._lib import load_library

        self.group = group if group is not None else dist.group.WORLD
        self.R = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        self.T, self.Hl, self.D, self.dtype = max_rows, heads_local, head_dim, dtype
        self.Hg = heads_local * self.R
        dev = torch.device(device)
        lib = load_library()
        self.ll_q = symm_mem.empty((lib.b200attn_peer_ll_bytes(max_rows, heads_local, head_dim, self.R, 0) // 8,), dtype=torch.int64, device=dev)
        self.ll_o = symm_mem.empty((lib.b200attn_peer_ll_bytes(max_rows, heads_local, head_dim, self.R, 1) // 8,), dtype=torch.int64, device=dev)
        self.ll_q.zero_(); self.ll_o.zero_()
        self.h_q = symm_mem.rendezvous(self.ll_q, self.group)
        self.h_o = symm_mem.rendezvous(self.ll_o, self.group)
        self.epoch = torch.zeros(2, 2, dtype=torch.int32, device=dev)
        self.gathered = torch.zeros(2, max_rows, self.Hg, head_dim, dtype=dtype, device=dev)
        self.o_part = torch.zeros(max_rows, self.Hg, head_dim, dtype=dtype, device=dev)
        self.lse_part = torch.full((max_rows, self.Hg), float("-inf"), dtype=torch.float32, device=dev)
        self.slot = 0
        torch.cuda.synchronize(dev)
        dist.barrier(self.group)      # every rank's receive words are zero before anyone can push

    def _dt(self):
        from ._lib import DT_BF16, DT_FP16

        return DT_BF16 if self.dtype == torch.bfloat16 else DT_FP16

    def gather_q(self, q_local: torch.Tensor) -> torch.Tensor:
        """q_local [n, Hl, D] (any token / head stride) -> the group's query heads [n, R * Hl, D] in a local buffer."""
        from ._lib import check, load_library

        n = q_local.shape[0]
        if n > self.T or q_local.shape[1] != self.Hl or q_local.shape[2] != self.D or q_local.stride(2) != 1 or q_local.dtype != self.dtype:
            raise ValueError("q_local must be [n <= max_rows, heads_local, head_dim] of the exchange dtype")
        self.slot ^= 1     # two local result buffers, so the caller may still read the previous step's queries
        out = self.gathered[self.slot]
        stream = torch.cuda.current_stream(q_local.device)
        check(load_library().b200attn_peer_gather_heads(
            q_local.data_ptr(), q_local.stride(0), q_local.stride(1), n, self.T, self.Hl, self.D, self._dt(), out.data_ptr(),
            self.h_q.buffer_ptrs_dev, self.epoch[0].data_ptr(), self.rank, self.R, stream.cuda_stream), "b200attn_peer_gather_heads")
        return out[:n]

    def partial_buffers(self, n: int) -> tuple[torch.Tensor, torch.Tensor]:
        return self.o_part[:n], self.lse_part[:n]

    def combine(self, n: int, out: torch.Tensor | None = None) -> torch.Tensor:
        from ._lib import check, load_library

        dev = self.o_part.device
        if out is None:
            out = torch.empty((n, self.Hl, self.D), dtype=self.dtype, device=dev)
        stream = torch.cuda.current_stream(dev)
        check(load_library().b200attn_peer_combine(
            self.o_part.data_ptr(), self.lse_part.data_ptr(), n, self.T, self.Hl, self.D, self._dt(), self.h_o.buffer_ptrs_dev,
            self.epoch[1].data_ptr(), self.rank, self.R, out.data_ptr(), out.stride(0), out.stride(1), stream.cuda_stream),
            "b200attn_peer_combine")
        return out
