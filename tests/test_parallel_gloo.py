"""world_size-2 gloo tests (CPU) of the multi-GPU host logic in ht_vllm_omni_b200/parallel.py:
head sharding + output all-gather, and the sequence-split (DCP) exchange.  The attention arithmetic of each rank is
stood in for by the numpy ORACLE here (tests may use it; the product's merge is the CUDA op and is exercised in the
-m gpu tests) — what these tests pin is the partitioning, the collectives and the recombination."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ht_vllm_omni_b200 import parallel as P
from oracle import oracle as O


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _spawn(fn, world=2, *args):
    port = _free_port()
    mp.spawn(_entry, args=(world, port, fn, args), nprocs=world, join=True)


def _entry(rank, world, port, fn, args):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        fn(rank, world, *args)
    finally:
        dist.destroy_process_group()


def _problem(seed=0, lens=(37, 200, 1), Hq=8, Hkv=2, D=128, bs=16):
    rng = np.random.default_rng(seed)
    lens = np.asarray(lens, np.int32)
    bt, nb = O.make_block_table(lens, bs, seed=99)
    kc = rng.standard_normal((nb, bs, Hkv, D)).astype(np.float32)
    vc = rng.standard_normal((nb, bs, Hkv, D)).astype(np.float32)
    kc, vc = O.bf16_round(kc), O.bf16_round(vc)
    q = O.bf16_round(rng.standard_normal((len(lens), Hq, D)).astype(np.float32))
    return lens, bt, kc, vc, q


def _oracle_merge(parts_o, parts_lse, out, lse_out):
    o, _ = O.merge_lse_ref(parts_o.float().numpy(), parts_lse.numpy())
    out.copy_(torch.from_numpy(o).to(out.dtype))


def test_head_shard_geometry():
    s = P.HeadShard(2, 1, 32, 4)
    assert (s.num_q_heads, s.num_kv_heads, s.replicas) == (16, 2, 1)
    assert list(s.q_head_range) == list(range(16, 32)) and list(s.kv_head_range) == [2, 3]
    s = P.HeadShard(8, 5, 32, 4)                      # tp > Hkv: KV head 2 replicated on ranks 4 and 5
    assert (s.num_q_heads, s.num_kv_heads, s.replicas, s.replica_rank) == (4, 1, 2, 1)
    assert list(s.kv_head_range) == [2] and s.replica_group_ranks() == [4, 5]
    assert all(q // 8 == 2 for q in s.q_head_range)   # its query heads all belong to that KV head
    for bad in ((3, 0, 32, 4), (8, 0, 30, 4), (2, 2, 32, 4)):
        with pytest.raises(ValueError):
            P.HeadShard(*bad)


def test_dcp_local_seq_lens_matches_positions():
    lens = torch.tensor([0, 1, 5, 16, 37, 1000], dtype=torch.int32)
    for dcp, il in ((2, 1), (4, 1), (2, 16), (3, 8)):
        tot = torch.zeros_like(lens)
        for r in range(dcp):
            got = P.dcp_local_seq_lens(lens, dcp, r, il)
            want = [len(P.dcp_local_positions(int(L), dcp, r, il)) for L in lens]
            assert got.tolist() == want, (dcp, il, r)
            tot += got
        assert tot.tolist() == lens.tolist()


def _head_sharded(rank, world):
    lens, bt, kc, vc, q = _problem()
    Hq, Hkv = q.shape[1], kc.shape[2]
    full = O.paged_decode_exact(q, kc, vc, bt, lens, 128 ** -0.5)
    sh = P.HeadShard(world, rank, Hq, Hkv)
    qs, ks = list(sh.q_head_range), list(sh.kv_head_range)
    local = O.paged_decode_exact(q[:, qs], kc[:, :, ks], vc[:, :, ks], bt, lens, 128 ** -0.5)   # own cache shard only
    gathered = P.all_gather_heads(torch.from_numpy(local.astype(np.float32)))
    assert gathered.shape == (len(lens), Hq, 128)
    np.testing.assert_allclose(gathered.numpy(), full, rtol=0, atol=1e-6)


def test_head_sharding_plus_all_gather_equals_unsharded():
    _spawn(_head_sharded, 2)


def _dcp(rank, world, interleave):
    # one KV head replicated on both ranks (tp = 2 > Hkv = 1); each rank owns Hq/2 query heads and every 2nd token run
    lens, bt, kc, vc, q = _problem(seed=3, lens=(37, 200, 1, 0), Hq=8, Hkv=1)
    full = O.paged_decode_exact(q, kc, vc, bt, lens, 128 ** -0.5)
    sh = P.HeadShard(world, rank, 8, 1)
    assert sh.replicas == world
    bs = kc.shape[1]
    # this rank's compacted cache: its positions in local order, in a fresh identity block table
    loc_lens = P.dcp_local_seq_lens(torch.from_numpy(lens), world, rank, interleave).numpy()
    nblk = max(1, int(-(-loc_lens.max() // bs)))
    lbt = np.arange(len(lens) * nblk, dtype=np.int32).reshape(len(lens), nblk)
    lk = np.zeros((len(lens) * nblk, bs, 1, 128), np.float32); lv = np.zeros_like(lk)
    for s, L in enumerate(lens):
        for j, p in enumerate(P.dcp_local_positions(int(L), world, rank, interleave)):
            src = (bt[s, p // bs], p % bs)
            lk[lbt[s, j // bs], j % bs] = kc[src]; lv[lbt[s, j // bs], j % bs] = vc[src]
    q_local = torch.from_numpy(q[:, list(sh.q_head_range)])
    q_group = P.dcp_gather_queries(q_local).numpy()             # all 8 heads of the group
    np.testing.assert_array_equal(q_group, q)
    # local attention over the slice + its LSE (fp64 definition)
    o = np.zeros_like(q_group); lse = np.full(q_group.shape[:2], -np.inf, np.float32)
    for s in range(len(lens)):
        n = int(loc_lens[s])
        if n == 0:
            continue
        K = np.stack([lk[lbt[s, j // bs], j % bs, 0] for j in range(n)]).astype(np.float64)
        V = np.stack([lv[lbt[s, j // bs], j % bs, 0] for j in range(n)]).astype(np.float64)
        logits = q_group[s].astype(np.float64) @ K.T * 128 ** -0.5
        m = logits.max(1, keepdims=True); e = np.exp(logits - m)
        o[s] = (e @ V) / e.sum(1, keepdims=True); lse[s] = (m[:, 0] + np.log(e.sum(1))).astype(np.float32)
    mine = P.dcp_combine(torch.from_numpy(o), torch.from_numpy(lse), merge_fn=_oracle_merge)
    valid = lens > 0
    np.testing.assert_allclose(mine.numpy()[valid], full[valid][:, list(sh.q_head_range)], rtol=0, atol=2e-5)
    with pytest.raises(Exception):    # default merge is the CUDA kernel: CPU tensors must fail loudly, not fall back
        P.dcp_combine(torch.from_numpy(o), torch.from_numpy(lse))


@pytest.mark.parametrize("interleave", [1, 16])
def test_sequence_split_exchange_equals_unsharded(interleave):
    _spawn(_dcp, 2, interleave)
