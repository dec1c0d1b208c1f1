"""CPU: the C-ABI library builds for sm_100a, loads without a GPU, and exports exactly what include/b200attn.h declares.
No compute entry point is exercised for results here (there is no GPU in the dev container); we only check that
argument validation returns error codes + messages instead of throwing, and that nothing silently succeeds."""
import ctypes as C
import os
import re
import subprocess

import pytest

from ht_vllm_omni_b200 import _build, _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b200attn.h")


@pytest.fixture(scope="module")
def lib():
    _build.build()
    return _lib.load_library()


def _declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"B200ATTN_API\s+[\w\s\*]+?\b(b200attn_\w+)\s*\(", src)))


def test_header_declares_the_expected_entry_points():
    syms = _declared_symbols()
    for name in ("b200attn_kv_write", "b200attn_decode_plan", "b200attn_decode", "b200attn_merge", "b200attn_merge_lse",
                 "b200attn_workspace_bytes", "b200attn_plan_bytes", "b200attn_last_error"):
        assert name in syms


def test_library_exports_every_declared_symbol(lib):
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.lib_path()], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r"\sT\s+(b200attn_\w+)", out))
    declared = set(_declared_symbols())
    assert declared <= exported, f"declared but not exported: {sorted(declared - exported)}"
    # nothing but the ABI (and the debug hook) leaks out of the library
    extra = exported - declared - {"b200attn_debug_set_timeline_buffer"}
    assert not extra, f"undeclared exports: {sorted(extra)}"


def test_ctypes_prototypes_cover_the_header(lib):
    declared = set(_declared_symbols())
    assert declared == set(_lib.PROTOTYPES), (declared ^ set(_lib.PROTOTYPES))
    # argument counts agree with the header text
    src = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    for name, (_, argtypes) in _lib.PROTOTYPES.items():
        m = re.search(r"\b%s\s*\(([^;]*?)\)\s*;" % name, src, flags=re.S)
        assert m, name
        args = m.group(1).strip()
        n = 0 if args in ("", "void") else len([a for a in args.split(",") if a.strip()])
        assert n == len(argtypes), f"{name}: header has {n} args, ctypes prototype {len(argtypes)}"


def test_only_sm100a_code_is_embedded():
    out = subprocess.run(["cuobjdump", "--list-elf", _lib.lib_path()], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    archs = set(re.findall(r"sm_(\d+a?)", out.stdout))
    assert archs == {"100a"}, archs


def test_version_and_geometry(lib):
    assert lib.b200attn_abi_version() == _lib.ABI_VERSION == 8
    assert _lib.TILE_TOKENS == 64 and _lib.MAX_GROUP == 16
    assert lib.b200attn_plan_bytes(64) >= (8 + 1024 * 64 + 1024 * 4 + 65) * 4   # header, 256-byte CTA entries, merge counters, prefix
    assert lib.b200attn_plan_bytes(64) % 16 == 0
    # workspace = [ctas][2][16][D] fp32 partial outputs + [ctas][2][16][2] fp32 (m, l)
    assert lib.b200attn_workspace_bytes(128, 296) == (296 * 2 * 16 * 128 + 296 * 2 * 16 * 2) * 4


def test_argument_errors_are_codes_not_exceptions(lib):
    assert lib.b200attn_set_option(b"no-such-option", 1) == -1
    assert b"unknown option" in lib.b200attn_last_error()
    # null pointers -> EINVAL with a message (validated before any CUDA call, so this works without a device)
    rc = lib.b200attn_kv_write(None, None, 0, 0, 0, 0, None, None, 0, 0, 0, None, 4, 2, 128, 16, 0, 2, None, None, 0, None)
    assert rc == -1 and b"null pointer" in lib.b200attn_last_error()
    rc = lib.b200attn_decode_plan(None, 4, 16, 2, 8, None, 0, None, 16, None, 0, None)
    assert rc == -1
    buf = (C.c_int32 * 16)()
    rc = lib.b200attn_decode_plan(buf, 4, 16, 3, 8, buf, 4, None, 16, buf, 64, None)  # Hq % Hkv != 0
    assert rc == -1 and b"bad shape" in lib.b200attn_last_error()
    rc = lib.b200attn_decode_plan(buf, 4, 16, 2, 8, buf, 4, None, 24, buf, 64, None)  # block_size not a multiple of 16
    assert rc == -1 and b"block_size" in lib.b200attn_last_error()
    rc = lib.b200attn_decode_plan(buf, 4, 16, 2, 8, buf, 4, None, 16, buf, 64, None)  # plan buffer too small
    assert rc == -3
    # empty batches are a successful no-op everywhere
    assert lib.b200attn_kv_write(None, None, 0, 0, 0, 0, None, None, 0, 0, 0, None, 0, 2, 128, 16, 0, 2, None, None, 0, None) == 0
    assert lib.b200attn_merge_lse(None, None, 0, 0, 0, None, None, 0, 16, 128, 0, None) == 0


def test_no_cpu_fallback_without_device(lib):
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present: the no-device behaviour is checked in the CPU container")
    assert lib.b200attn_sm_count() == 0
    assert lib.b200attn_default_ctas(2) == 0
    import ht_vllm_omni_b200.ops as ops

    with pytest.raises(_lib.B200AttnError):
        ops.default_ctas(torch.uint8)
    k = torch.zeros(4, 2, 128, dtype=torch.bfloat16)
    cache = torch.zeros(2, 16, 2, 128, dtype=torch.uint8)
    one = torch.ones(1)
    with pytest.raises((_lib.B200AttnError, NotImplementedError, RuntimeError)):
        torch.ops.b200attn.kv_write(k, k, cache, cache.clone(), torch.zeros(4, dtype=torch.int64), one, one)
    assert int(cache.sum()) == 0  # nothing was written on the CPU


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "ht_vllm_omni_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f
                assert "liboracle" not in text, f


def test_decode_static_refuses_oversize_requests_before_touching_the_device(lib):
    """b200attn_decode_static validates against the static array's limits first: no device is needed to be told no."""
    def call(num_seqs, head_dim):
        return lib.b200attn_decode_static(None, 0, 0, None, 0, 0, None, None, 0, 0, 0, 1, None, 0, None, num_seqs, 16, 2, head_dim, 16,
                                          _lib.DT_BF16, _lib.DT_FP8_E4M3, 1.0, None, None, 0, None, -1, None)

    assert call(1025, 128) == -3 and b"1024" in lib.b200attn_last_error()      # B200ATTN_ENOSPACE, names the limit
    assert call(8, 64) == -1 and b"head_dim" in lib.b200attn_last_error()      # B200ATTN_EINVAL
    assert call(0, 128) == 0                                                   # empty batch: nothing to do
