"""-m gpu: the vLLM plugin classes end to end on the device — MetadataBuilder.build -> Impl.do_kv_cache_update ->
Impl.forward — on the KV cache exactly as vLLM allocates it (raw int8 buffer -> get_kv_cache_shape /
get_kv_cache_stride_order), against the CPU oracle.  Dummy config objects stand in for VllmConfig, as the reference's
worker tests do."""
import types

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

vb = pytest.importorskip("vllm.v1.attention.backend")

from oracle import oracle as O  # noqa: E402


def _cfg(Hq, Hkv, D, max_tokens):
    mc = types.SimpleNamespace(get_num_attention_heads=lambda pc: Hq, get_num_kv_heads=lambda pc: Hkv, get_head_size=lambda: D)
    return types.SimpleNamespace(model_config=mc, parallel_config=None,
                                 scheduler_config=types.SimpleNamespace(max_num_batched_tokens=max_tokens))


def _alloc_like_vllm(be, nb, bs, Hkv, D, kv_cache_dtype, torch_dtype):
    shape = be.get_kv_cache_shape(nb, bs, Hkv, D, kv_cache_dtype)
    order = be.get_kv_cache_stride_order()
    phys = tuple(shape[i] for i in order)
    inv = [order.index(i) for i in range(len(order))]
    nbytes = int(np.prod(shape)) * torch.empty(0, dtype=torch_dtype).element_size()
    raw = torch.zeros(nbytes, dtype=torch.int8, device="cuda")
    return raw.view(torch_dtype).view(phys).permute(*inv)


@pytest.mark.parametrize("contract", ["forward_writes_kv", "separate_kv_update"])
@pytest.mark.parametrize("kv_cache_dtype", ["fp8", "auto"])
@pytest.mark.parametrize("q_lens", [[1, 1, 1, 1], [1, 3, 2, 7]], ids=["pure_decode", "mixed_multi_token_rows"])
def test_backend_step_matches_oracle(kv_cache_dtype, q_lens, contract, monkeypatch):
    """Both contracts vLLM knows (attention.py:493): forward() owns the KV write (the default: on a pure-decode step the
    store happens inside the decode launch), or unified_kv_cache_update -> do_kv_cache_update precedes forward()."""
    from ht_vllm_omni_b200 import backend as B

    be = B.B200PagedAttentionBackend
    monkeypatch.setattr(be, "forward_includes_kv_cache_update", contract == "forward_writes_kv")
    Hq, Hkv, D, bs = 16, 2, 128, 16
    fp8 = kv_cache_dtype == "fp8"
    ctx = np.array([300, 40, 0, 1000], np.int32)              # tokens already in the cache
    q_lens = np.array(q_lens, np.int32)
    seq = ctx + q_lens
    B_, T = len(seq), int(q_lens.sum())
    bt_np, nb = O.make_block_table(seq, bs, seed=5)
    g = torch.Generator().manual_seed(11)
    rnd = lambda *s: torch.randn(*s, generator=g).to(torch.bfloat16)  # noqa: E731
    # history: all tokens of every sequence (the last q_len of each are written by THIS step through the hook)
    hist_k = [rnd(int(L), Hkv, D) for L in seq]
    hist_v = [rnd(int(L), Hkv, D) for L in seq]
    k_scale = torch.tensor([0.02, 0.03]) if fp8 else torch.ones(1)
    v_scale = torch.tensor([0.025, 0.015]) if fp8 else torch.ones(1)
    # ---- oracle cache: everything written through the oracle's write ----
    shape = (nb, bs, Hkv, D)
    kc_o = np.zeros(shape, np.uint8 if fp8 else np.float32); vc_o = np.zeros_like(kc_o)
    pos_all = [(r, p) for r in range(B_) for p in range(int(seq[r]))]
    O.kv_write_ref(torch.cat(hist_k).float().numpy(), torch.cat(hist_v).float().numpy(), kc_o, vc_o,
                   O.slots_for_positions(bt_np, pos_all, bs), k_scale.numpy(), v_scale.numpy(), fp8=fp8)
    # ---- device: vLLM-style allocation, context pre-loaded, this step's tokens go through do_kv_cache_update ----
    cache_dt = torch.uint8 if fp8 else torch.bfloat16
    kv_cache = _alloc_like_vllm(be, nb, bs, Hkv, D, kv_cache_dtype, cache_dt)
    layer = types.SimpleNamespace(_k_scale=k_scale.cuda(), _v_scale=v_scale.cuda(), _k_scale_float=1.0, _v_scale_float=1.0)
    impl = be.get_impl_cls()(Hq, D, D ** -0.5, Hkv, None, None, kv_cache_dtype)
    pos_ctx = [(r, p) for r in range(B_) for p in range(int(ctx[r]))]
    if pos_ctx:
        ck = torch.cat([hist_k[r][: ctx[r]] for r in range(B_)]).reshape(-1, Hkv * D).cuda()
        cv = torch.cat([hist_v[r][: ctx[r]] for r in range(B_)]).reshape(-1, Hkv * D).cuda()
        impl.do_kv_cache_update(layer, ck, cv, kv_cache, torch.from_numpy(O.slots_for_positions(bt_np, pos_ctx, bs)).cuda())
    # this step: fused qkv activation (strided q/k/v slices), padded slot_mapping (-1) as under CUDA graphs
    pad = 3
    qkv = torch.zeros(T + pad, (Hq + 2 * Hkv) * D, dtype=torch.bfloat16)
    new_k = torch.cat([hist_k[r][ctx[r]:] for r in range(B_)]); new_v = torch.cat([hist_v[r][ctx[r]:] for r in range(B_)])
    q = rnd(T, Hq, D)
    qkv[:T, : Hq * D] = q.reshape(T, -1); qkv[:T, Hq * D:(Hq + Hkv) * D] = new_k.reshape(T, -1); qkv[:T, (Hq + Hkv) * D:] = new_v.reshape(T, -1)
    qkv = qkv.cuda()
    d_q, d_k, d_v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
    pos_new = [(r, p) for r in range(B_) for p in range(int(ctx[r]), int(seq[r]))]
    slots = np.concatenate([O.slots_for_positions(bt_np, pos_new, bs), -np.ones(pad, np.int64)])
    qsl = np.concatenate([[0], np.cumsum(q_lens)]).astype(np.int32)
    cm = vb.CommonAttentionMetadata(
        query_start_loc=torch.from_numpy(qsl).cuda(), query_start_loc_cpu=torch.from_numpy(qsl),
        seq_lens=torch.from_numpy(seq).cuda(), num_reqs=B_, num_actual_tokens=T, max_query_len=int(q_lens.max()),
        max_seq_len=int(seq.max()), block_table_tensor=torch.from_numpy(bt_np).cuda(), slot_mapping=torch.from_numpy(slots).cuda())
    spec = types.SimpleNamespace(block_size=bs, dtype=cache_dt)
    builder = be.get_builder_cls()(spec, ["layer.0"], _cfg(Hq, Hkv, D, 64), torch.device("cuda"))
    md = builder.build(0, cm)
    if not be.forward_includes_kv_cache_update:
        impl.do_kv_cache_update(layer, d_k, d_v, kv_cache, cm.slot_mapping)
    output = torch.full((T + pad, Hq * D), 7.0, dtype=torch.bfloat16, device="cuda")
    ret = impl.forward(layer, d_q, d_k, d_v, kv_cache, md, output)
    torch.cuda.synchronize()
    assert ret is output and bool((output[T:] == 7.0).all()), "rows past num_actual_tokens must be left alone"
    # cache bytes bit-exact vs the oracle (logical [nb, 2, bs, Hkv, D] indexing, as the KV hand-off code reads it)
    got_k = kv_cache[:, 0].contiguous().cpu(); got_v = kv_cache[:, 1].contiguous().cpu()
    if fp8:
        assert np.array_equal(got_k.numpy(), kc_o) and np.array_equal(got_v.numpy(), vc_o)
    else:
        assert np.array_equal(got_k.float().numpy(), kc_o) and np.array_equal(got_v.float().numpy(), vc_o)
    # attention: token t of request r sees ctx[r] + i + 1 positions
    row_req = np.repeat(np.arange(B_), q_lens)
    row_len = np.concatenate([ctx[r] + 1 + np.arange(q_lens[r]) for r in range(B_)]).astype(np.int32)
    ref = O.paged_decode_exact(q.float().numpy(), kc_o, vc_o, bt_np[row_req], row_len, D ** -0.5, k_scale.numpy(), v_scale.numpy(), fp8=fp8)
    got = output[:T].view(T, Hq, D).float().cpu().numpy()
    tol = 1e-2 if fp8 else 2e-3
    err = np.abs(got - ref)
    assert np.all(err <= tol + np.maximum(np.abs(ref), 2.0 ** -14) * 2.0 ** -7), float(err.max())


@pytest.mark.parametrize("contract", ["forward_writes_kv", "separate_kv_update"])
def test_backend_under_cuda_graph_replay_with_changing_seq_lens(contract, monkeypatch):
    """FULL-graph decode: capture once at a padded batch, replay after the runner updates seq_lens / block_table /
    slot_mapping IN PLACE (the plan kernel re-reads them on the device; nothing is baked in at capture).  Under the
    forward-writes contract the captured launch is the decode that also stores the tokens (slots -1 here: padding)."""
    from ht_vllm_omni_b200 import backend as B

    be = B.B200PagedAttentionBackend
    monkeypatch.setattr(be, "forward_includes_kv_cache_update", contract == "forward_writes_kv")
    Hq, Hkv, D, bs, Bp = 16, 2, 128, 16, 8
    impl = be.get_impl_cls()(Hq, D, D ** -0.5, Hkv, None, None, "fp8")
    layer = types.SimpleNamespace(_k_scale=torch.tensor([0.02], device="cuda"), _v_scale=torch.tensor([0.02], device="cuda"))
    lens_a = np.array([100, 2000, 37, 0, 0, 0, 0, 0], np.int32)      # 3 live rows + padding (seq_len 0)
    lens_b = np.array([101, 2001, 38, 512, 64, 0, 0, 0], np.int32)
    bt_np, nb = O.make_block_table(np.maximum(lens_a, lens_b), bs, seed=3)
    kv_cache = _alloc_like_vllm(be, nb, bs, Hkv, D, "fp8", torch.uint8)
    rng = np.random.default_rng(0)
    kc_o = rng.integers(0, 0x78, (nb, bs, Hkv, D), dtype=np.uint8); vc_o = rng.integers(0, 0x78, (nb, bs, Hkv, D), dtype=np.uint8)
    kv_cache[:, 0].copy_(torch.from_numpy(kc_o)); kv_cache[:, 1].copy_(torch.from_numpy(vc_o))
    d_seq = torch.from_numpy(lens_a).cuda(); d_bt = torch.from_numpy(bt_np).cuda()
    d_slots = torch.full((Bp,), -1, dtype=torch.int64, device="cuda")
    qsl = torch.arange(Bp + 1, dtype=torch.int32)
    q = torch.randn(Bp, Hq * D, generator=torch.Generator().manual_seed(1)).to(torch.bfloat16).cuda()
    kvn = torch.zeros(Bp, Hkv * D, dtype=torch.bfloat16, device="cuda")
    out = torch.zeros(Bp, Hq * D, dtype=torch.bfloat16, device="cuda")
    cm = vb.CommonAttentionMetadata(query_start_loc=qsl.cuda(), query_start_loc_cpu=qsl, seq_lens=d_seq, num_reqs=Bp,
                                    num_actual_tokens=Bp, max_query_len=1, max_seq_len=4096, block_table_tensor=d_bt, slot_mapping=d_slots)
    builder = be.get_builder_cls()(types.SimpleNamespace(block_size=bs, dtype=torch.uint8), ["l0"], _cfg(Hq, Hkv, D, 64), torch.device("cuda"))

    def step():
        md = builder.build_for_cudagraph_capture(cm)
        if not be.forward_includes_kv_cache_update:
            impl.do_kv_cache_update(layer, kvn, kvn, kv_cache, d_slots)     # all slots -1: writes nothing
        impl.forward(layer, q, kvn, kvn, kv_cache, md, out)

    s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        step()
    torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        step()
    for lens in (lens_a, lens_b, lens_a):
        d_seq.copy_(torch.from_numpy(lens))
        out.zero_()
        graph.replay()
        torch.cuda.synchronize()
        ref = O.paged_decode_exact(q.view(Bp, Hq, D).float().cpu().numpy(), kc_o, vc_o, bt_np, lens, D ** -0.5,
                                   np.array([0.02], np.float32), np.array([0.02], np.float32), fp8=True)
        got = out.view(Bp, Hq, D).float().cpu().numpy()
        live = lens > 0
        err = np.abs(got[live] - ref[live])
        assert np.all(err <= 1e-2 + np.maximum(np.abs(ref[live]), 2.0 ** -14) * 2.0 ** -7), float(err.max())


def test_full_graph_captures_forward_only_and_build_runs_eagerly_before_each_replay(monkeypatch):
    """Two-op contract (B200ATTN_SEPARATE_KV_UPDATE=1), the one under which the builder declares AttentionCGSupport.ALWAYS.
    vLLM's FULL-cudagraph flow: `build()` runs eagerly before every replay, only the layer hooks are inside
    the graph.  One graph per token count serves BOTH a batch with multi-token requests and a pure-decode batch of the same
    size, so the captured kernels must read their row lengths / row map from buffers that do not depend on which of the two
    the batch is.  Capture on a mixed batch (3 requests, q_lens 2+1+1), replay on a pure-decode batch (4 requests), then on
    the mixed batch again."""
    from ht_vllm_omni_b200 import backend as B

    be = B.B200PagedAttentionBackend
    monkeypatch.setattr(be, "forward_includes_kv_cache_update", False)
    assert be.get_builder_cls().get_cudagraph_support(None, None) == vb.AttentionCGSupport.ALWAYS
    Hq, Hkv, D, bs, T, R = 16, 2, 128, 16, 4, 4
    impl = be.get_impl_cls()(Hq, D, D ** -0.5, Hkv, None, None, "fp8")
    layer = types.SimpleNamespace(_k_scale=torch.tensor([0.02], device="cuda"), _v_scale=torch.tensor([0.02], device="cuda"))
    max_lens = np.array([700, 300, 1500, 64], np.int32)
    bt_np, nb = O.make_block_table(max_lens, bs, seed=4)
    kv_cache = _alloc_like_vllm(be, nb, bs, Hkv, D, "fp8", torch.uint8)
    rng = np.random.default_rng(1)
    kc_o = rng.integers(0, 0x78, (nb, bs, Hkv, D), dtype=np.uint8); vc_o = rng.integers(0, 0x78, (nb, bs, Hkv, D), dtype=np.uint8)
    kv_cache[:, 0].copy_(torch.from_numpy(kc_o)); kv_cache[:, 1].copy_(torch.from_numpy(vc_o))
    # persistent runner buffers (capture and replay see the same addresses)
    d_seq = torch.zeros(R, dtype=torch.int32, device="cuda"); d_bt = torch.from_numpy(bt_np).cuda()
    d_qsl = torch.zeros(R + 1, dtype=torch.int32, device="cuda")
    d_slots = torch.full((T,), -1, dtype=torch.int64, device="cuda")
    q = torch.randn(T, Hq * D, generator=torch.Generator().manual_seed(2)).to(torch.bfloat16).cuda()
    kvn = torch.zeros(T, Hkv * D, dtype=torch.bfloat16, device="cuda")
    out = torch.zeros(T, Hq * D, dtype=torch.bfloat16, device="cuda")
    builder = be.get_builder_cls()(types.SimpleNamespace(block_size=bs, dtype=torch.uint8), ["l0"], _cfg(Hq, Hkv, D, 64), torch.device("cuda"))

    def metadata(seq, q_lens):
        n = len(seq)
        qsl = np.concatenate([[0], np.cumsum(q_lens)]).astype(np.int32)
        d_seq[:n].copy_(torch.from_numpy(np.asarray(seq, np.int32))); d_qsl[:n + 1].copy_(torch.from_numpy(qsl))
        return vb.CommonAttentionMetadata(query_start_loc=d_qsl[:n + 1], query_start_loc_cpu=torch.from_numpy(qsl), seq_lens=d_seq[:n],
                                          num_reqs=n, num_actual_tokens=T, max_query_len=int(max(q_lens)), max_seq_len=2048,
                                          block_table_tensor=d_bt[:n], slot_mapping=d_slots)

    def reference(seq, q_lens):
        row_req = np.repeat(np.arange(len(seq)), q_lens)
        row_len = np.concatenate([seq[r] - q_lens[r] + 1 + np.arange(q_lens[r]) for r in range(len(seq))]).astype(np.int32)
        return O.paged_decode_exact(q.view(T, Hq, D).float().cpu().numpy(), kc_o, vc_o, bt_np[row_req], row_len, D ** -0.5,
                                    np.array([0.02], np.float32), np.array([0.02], np.float32), fp8=True)

    mixed = (np.array([700, 300, 1500]), np.array([2, 1, 1]))
    pure = (np.array([650, 280, 1400, 64]), np.array([1, 1, 1, 1]))
    md = builder.build(0, metadata(*mixed))

    def hooks():
        impl.do_kv_cache_update(layer, kvn, kvn, kv_cache, d_slots)     # all slots -1: writes nothing
        impl.forward(layer, q, kvn, kvn, kv_cache, md, out)

    s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        hooks()
    torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        hooks()
    for seq, q_lens in (pure, mixed, pure):
        builder.build(0, metadata(seq, q_lens))     # eager, outside the graph -- as vLLM does
        out.zero_()
        graph.replay()
        torch.cuda.synchronize()
        ref = reference(seq, q_lens)
        got = out.view(T, Hq, D).float().cpu().numpy()
        err = np.abs(got - ref)
        assert np.all(err <= 1e-2 + np.maximum(np.abs(ref), 2.0 ** -14) * 2.0 ** -7), float(err.max())


def test_uniform_spec_decode_batch_under_full_graph_with_the_forward_writes_contract():
    """AttentionCGSupport.UNIFORM_BATCH under the default contract: a FULL graph captured for a uniform batch of q_len = 1 + k
    (spec-decode verification / the talker's MTP step) holds forward() only -- KV write of the 1 + k tokens, the rows, the
    prefill kernel -- and is replayed for other sequence lengths and for a padded batch (fewer live requests); build() runs
    eagerly before every replay, as in vLLM."""
    from ht_vllm_omni_b200 import backend as B

    be = B.B200PagedAttentionBackend
    assert be.forward_includes_kv_cache_update
    assert be.get_builder_cls().get_cudagraph_support(None, None) == vb.AttentionCGSupport.UNIFORM_BATCH
    Hq, Hkv, D, bs, R, QL = 16, 2, 128, 16, 4, 3
    T = R * QL
    impl = be.get_impl_cls()(Hq, D, D ** -0.5, Hkv, None, None, "fp8")
    ks, vs = torch.tensor([0.02], device="cuda"), torch.tensor([0.03], device="cuda")
    layer = types.SimpleNamespace(_k_scale=ks, _v_scale=vs)
    max_lens = np.array([700, 300, 1500, 64], np.int32)
    bt_np, nb = O.make_block_table(max_lens, bs, seed=4)
    kv_cache = _alloc_like_vllm(be, nb, bs, Hkv, D, "fp8", torch.uint8)
    rng = np.random.default_rng(1)
    base_k = rng.integers(0, 0x78, (nb, bs, Hkv, D), dtype=np.uint8); base_v = rng.integers(0, 0x78, (nb, bs, Hkv, D), dtype=np.uint8)
    d_seq = torch.zeros(R, dtype=torch.int32, device="cuda"); d_bt = torch.from_numpy(bt_np).cuda()
    d_qsl = torch.zeros(R + 1, dtype=torch.int32, device="cuda")
    d_slots = torch.full((T,), -1, dtype=torch.int64, device="cuda")
    g = torch.Generator().manual_seed(2)
    q = torch.randn(T, Hq * D, generator=g).to(torch.bfloat16).cuda()
    kn = torch.randn(T, Hkv * D, generator=g).to(torch.bfloat16).cuda(); vn = torch.randn(T, Hkv * D, generator=g).to(torch.bfloat16).cuda()
    out = torch.zeros(T, Hq * D, dtype=torch.bfloat16, device="cuda")
    builder = be.get_builder_cls()(types.SimpleNamespace(block_size=bs, dtype=torch.uint8), ["l0"], _cfg(Hq, Hkv, D, 64), torch.device("cuda"))

    def metadata(seq, live):
        """`live` requests of QL tokens each, the rest of the captured batch is padding (seq_len 0, repeated offsets, slot -1)"""
        qsl = np.minimum(np.arange(R + 1), live) * QL
        sl = np.zeros(R, np.int32); sl[:live] = seq[:live]
        d_seq.copy_(torch.from_numpy(sl)); d_qsl.copy_(torch.from_numpy(qsl.astype(np.int32)))
        pos = [(r, p) for r in range(live) for p in range(int(seq[r]) - QL, int(seq[r]))]
        slots = np.full(T, -1, np.int64); slots[:live * QL] = O.slots_for_positions(bt_np, pos, bs)
        d_slots.copy_(torch.from_numpy(slots))
        return vb.CommonAttentionMetadata(query_start_loc=d_qsl, query_start_loc_cpu=torch.from_numpy(qsl.astype(np.int32)), seq_lens=d_seq,
                                          num_reqs=R, num_actual_tokens=T, max_query_len=QL, max_seq_len=2048, block_table_tensor=d_bt,
                                          slot_mapping=d_slots), slots

    def reference(seq, live, slots):
        kc, vc = base_k.copy(), base_v.copy()
        O.kv_write_ref(kn.view(T, Hkv, D).float().cpu().numpy(), vn.view(T, Hkv, D).float().cpu().numpy(), kc, vc, slots, np.array([0.02], np.float32),
                       np.array([0.03], np.float32), fp8=True)
        rows = live * QL
        row_req = np.repeat(np.arange(live), QL)
        row_len = np.concatenate([seq[r] - QL + 1 + np.arange(QL) for r in range(live)]).astype(np.int32)
        return O.paged_decode_exact(q.view(T, Hq, D).float().cpu().numpy()[:rows], kc, vc, bt_np[row_req], row_len, D ** -0.5,
                                    np.array([0.02], np.float32), np.array([0.03], np.float32), fp8=True), rows

    first = (np.array([700, 300, 1500, 64]), 4)
    cm, _ = metadata(*first)
    md = builder.build_for_cudagraph_capture(cm)
    assert not md.pure_decode

    def hooks():
        impl.forward(layer, q, kn, vn, kv_cache, md, out)

    s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        hooks()
    torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        hooks()
    for seq, live in (first, (np.array([650, 280, 1400, 40]), 4), (np.array([123, 999, 0, 0]), 2)):
        kv_cache[:, 0].copy_(torch.from_numpy(base_k)); kv_cache[:, 1].copy_(torch.from_numpy(base_v))
        cm, slots = metadata(seq, live)
        builder.build(0, cm)            # eager, outside the graph
        out.zero_()
        graph.replay()
        torch.cuda.synchronize()
        ref, rows = reference(seq, live, slots)
        got = out.view(T, Hq, D).float().cpu().numpy()[:rows]
        err = np.abs(got - ref)
        assert np.all(err <= 1e-2 + np.maximum(np.abs(ref), 2.0 ** -14) * 2.0 ** -7), (live, float(err.max()))
