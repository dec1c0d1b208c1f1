"""b200attn_decode_static (include/b200attn.h): the decode entry point that takes its plan buffer and split-KV workspace from a
file-scope device array inside the library instead of the caller, and allocates nothing.  It has no counterpart in the
reference, so the checker here is the existing entry point pair
b200attn_decode_plan + b200attn_decode on the same inputs, through the same C-ABI, within the tolerance the parity tests use
(BASELINE.json north_star: max-abs 1e-2 with an FP8 cache, 2e-3 with a 16-bit one).

Collected last on purpose (file name): a failure here must not hide other tests.
Runs alone in a fresh process and after the other tests alike: it owns every buffer it uses and sets no library option."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

D, BS = 128, 16
TOL = {"fp8": 1e-2, "bf16": 2e-3}


def _problem(lens, Hq, Hkv, kv, seed):
    g = torch.Generator().manual_seed(seed)
    lens = np.asarray(lens, np.int32)
    B = len(lens)
    nblk = max(1, -(-int(lens.max()) // BS))
    nb = B * nblk + 3
    bt = torch.randperm(nb, generator=g)[: B * nblk].to(torch.int32).reshape(B, nblk).cuda()
    if kv == "fp8":
        phys = torch.randint(0, 0x78, (nb, 2, Hkv, BS, D), generator=g, dtype=torch.uint8)
        phys = (phys | (torch.randint(0, 2, phys.shape, generator=g, dtype=torch.uint8) << 7)).cuda()   # random signs
        ks = torch.tensor([0.02], device="cuda"); vs = torch.tensor([0.03], device="cuda")
    else:
        phys = torch.randn((nb, 2, Hkv, BS, D), generator=g).to(torch.bfloat16).cuda()
        ks = torch.ones(1, device="cuda"); vs = torch.ones(1, device="cuda")
    kc, vc = phys.permute(0, 1, 3, 2, 4).unbind(1)          # logical [nb, bs, Hkv, D] over HND memory
    q = torch.randn(B, Hq, D, generator=g).to(torch.bfloat16).cuda()
    return dict(B=B, Hq=Hq, Hkv=Hkv, kv=kv, lens=torch.from_numpy(lens).cuda(), lens_np=lens, bt=bt, kc=kc, vc=vc, ks=ks, vs=vs, q=q,
                phys=phys)


def _common_args(lib_mod, P, out):
    kc, vc, q = P["kc"], P["vc"], P["q"]
    assert kc.stride() == vc.stride()
    kvdt = lib_mod.DT_FP8_E4M3 if P["kv"] == "fp8" else lib_mod.DT_BF16
    return [q.data_ptr(), q.stride(0), q.stride(1), out.data_ptr(), out.stride(0), out.stride(1), kc.data_ptr(), vc.data_ptr(),
            kc.stride(0), kc.stride(1), kc.stride(2), kc.shape[0], P["bt"].data_ptr(), P["bt"].stride(0), P["lens"].data_ptr(), P["B"],
            P["Hq"], P["Hkv"], D, BS, lib_mod.DT_BF16, kvdt, float(D ** -0.5), P["ks"].data_ptr(), P["vs"].data_ptr(), 0]


def _existing(lib, lib_mod, P, ctas):
    """b200attn_decode_plan + b200attn_decode with caller-owned plan / workspace (the entry points every other test covers)"""
    out = torch.full((P["B"], P["Hq"], D), float("nan"), dtype=torch.bfloat16, device="cuda")
    lse = torch.zeros(P["B"], P["Hq"], dtype=torch.float32, device="cuda")
    plan = torch.zeros(lib.b200attn_plan_bytes(P["B"]) // 4, dtype=torch.int32, device="cuda")
    ws = torch.empty(lib.b200attn_workspace_bytes(D, ctas) // 4, dtype=torch.float32, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    rc = lib.b200attn_decode_plan(P["lens"].data_ptr(), P["B"], P["Hq"], P["Hkv"], ctas, P["bt"].data_ptr(), P["bt"].stride(0), None, BS,
                                  plan.data_ptr(), plan.numel() * 4, stream)
    assert rc == 0, lib.b200attn_last_error()
    rc = lib.b200attn_decode(*_common_args(lib_mod, P, out), plan.data_ptr(), ctas, ws.data_ptr(), ws.numel() * 4, lse.data_ptr(), None,
                             -1, lib_mod.DECODE_FUSED_MERGE, stream)
    assert rc == 0, lib.b200attn_last_error()
    torch.cuda.synchronize()
    return out, lse


def _outputs(P):
    out = torch.full((P["B"], P["Hq"], D), float("nan"), dtype=torch.bfloat16, device="cuda")
    lse = torch.zeros(P["B"], P["Hq"], dtype=torch.float32, device="cuda")
    return out, lse


def _static(lib, lib_mod, P, outputs=None):
    out, lse = outputs if outputs is not None else _outputs(P)
    rc = lib.b200attn_decode_static(*_common_args(lib_mod, P, out), lse.data_ptr(), -1, torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.b200attn_last_error()
    torch.cuda.synchronize()
    return out, lse


def _check(got, ref, P):
    (o_s, l_s), (o_e, l_e) = got, ref
    live = torch.from_numpy(P["lens_np"] > 0).cuda()
    a, b = o_s[live].float(), o_e[live].float()
    assert torch.isfinite(a).all() and torch.isfinite(b).all()
    err = float((a - b).abs().max())
    assert err <= TOL[P["kv"]], f"static vs caller-owned scratch: max abs {err:.3e} > {TOL[P['kv']]}"
    assert float((l_s[live] - l_e[live]).abs().max()) <= 2e-3
    assert float(o_s[~live].float().abs().sum()) == 0 and bool(torch.isneginf(l_s[~live]).all())   # empty rows: zeros, lse = -inf
    return err


def test_decode_static_matches_the_caller_owned_scratch_entry_point_on_consecutive_calls():
    import ht_vllm_omni_b200._lib as lib_mod

    lib = lib_mod.load_library()          # raises when libb200attn.so is missing: no fallback
    ctas = min(lib.b200attn_default_ctas(lib_mod.DT_FP8_E4M3), 320)
    # call 1: ragged FP8 batch whose long rows are cut across many CTAs (in-kernel merge through the static workspace)
    P1 = _problem([1, 37, 5000, 64, 0, 20000, 513, 2048], 16, 2, "fp8", seed=11)
    # call 2, right after it on the same stream: other shapes and lengths, so whatever call 1 left in the array is stale
    P2 = _problem([9000, 3, 700, 0, 129], 32, 4, "fp8", seed=12)
    # call 3: 16-bit cache (the HMMA kernel) over the same array
    P3 = _problem([3000, 250, 17], 16, 2, "bf16", seed=13)
    errs = []
    first = _static(lib, lib_mod, P1)
    second = _static(lib, lib_mod, P2)     # two calls in a row, nothing in between
    errs.append(_check(first, _existing(lib, lib_mod, P1, ctas), P1))
    errs.append(_check(second, _existing(lib, lib_mod, P2, ctas), P2))
    # the device's free memory must not shrink across two calls (outputs are allocated BEFORE the reading, so torch's caching
    # allocator has no reason to ask the driver for memory in between)
    o3, o1 = _outputs(P3), _outputs(P1)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    third = _static(lib, lib_mod, P3, o3)
    again = _static(lib, lib_mod, P1, o1)  # and the first problem once more after the array has held two other plans
    assert torch.cuda.mem_get_info()[0] >= free_before, "the static entry point took device memory"
    errs.append(_check(third, _existing(lib, lib_mod, P3, ctas), P3))
    errs.append(_check(again, first, P1))
    print(f"[no-alloc] max abs vs b200attn_decode: {['%.2e' % e for e in errs]}")


def test_decode_static_rejects_what_the_static_array_cannot_hold():
    import ht_vllm_omni_b200._lib as lib_mod

    lib = lib_mod.load_library()
    P = _problem([100, 200], 16, 2, "fp8", seed=3)
    out = torch.zeros(P["B"], P["Hq"], D, dtype=torch.bfloat16, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    args = _common_args(lib_mod, P, out)
    too_many = list(args); too_many[15] = 1025          # num_seqs beyond B200ATTN_STATIC_MAX_SEQS: refused before any launch
    assert lib.b200attn_decode_static(*too_many, None, -1, stream) == -3   # B200ATTN_ENOSPACE
    assert b"1024" in lib.b200attn_last_error()
    other_d = list(args); other_d[18] = 64               # head_dim the static workspace is not laid out for
    assert lib.b200attn_decode_static(*other_d, None, -1, stream) == -1   # B200ATTN_EINVAL
    assert b"head_dim" in lib.b200attn_last_error()
    torch.cuda.synchronize()
    assert float(out.abs().sum()) == 0                   # neither call launched anything
    assert ctypes.sizeof(ctypes.c_void_p) == 8
