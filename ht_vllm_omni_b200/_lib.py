"""ctypes binding of libb200attn.so (include/b200attn.h).  No torch types cross this boundary.

The library is built in-tree by ``python -m ht_vllm_omni_b200._build`` (or ``__graft_entry__.build()``).
There is deliberately no fallback: if the shared object is missing or a call fails, we raise.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

DT_BF16, DT_FP16, DT_FP8_E4M3 = 0, 1, 2
TILE_TOKENS = 64
MAX_GROUP = 16
MAX_CTAS = 1024
PEER_CTAS = 32
ABI_VERSION = 8
DECODE_FUSED_MERGE = 1

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
_LOCK = threading.Lock()


class B200AttnError(RuntimeError):
    """A libb200attn entry point returned a non-zero code."""


def lib_path() -> str:
    return os.path.join(_HERE, "libb200attn.so")


_i32, _i64, _f32, _sz, _p = C.c_int32, C.c_int64, C.c_float, C.c_size_t, C.c_void_p

# name -> (restype, argtypes); mirrors include/b200attn.h one to one (tests/test_abi.py cross-checks).
PROTOTYPES = {
    "b200attn_abi_version": (C.c_int, []),
    "b200attn_last_error": (C.c_char_p, []),
    "b200attn_sm_count": (C.c_int, []),
    "b200attn_kv_write": (C.c_int, [_p, _p, _i64, _i64, _i64, _i64, _p, _p, _i64, _i64, _i64, _p, _i32, _i32, _i32, _i32,
                                    _i32, _i32, _p, _p, _i32, _p]),
    "b200attn_kv_write_per_token_head": (C.c_int, [_p, _p, _i64, _i64, _i64, _i64, _p, _p, _i64, _i64, _i64, _p, _p, _i64, _i64, _i64,
                                                   _p, _i32, _i32, _i32, _i32, _i32, _i32, _p]),
    "b200attn_rope_kv_write": (C.c_int, [_p, _i64, _i64, _i32, _p, _p, _i64, _i64, _i64, _i64, _p, _p, _i32, _i32, _p, _p,
                                         _i64, _i64, _i64, _p, _i32, _i32, _i32, _i32, _i32, _i32, _p, _p, _i32, _p]),
    "b200attn_norm_rope_kv_write": (C.c_int, [_p, _i64, _i64, _i32, _p, _p, _i64, _i64, _i64, _i64, _p, _i64, _i32, _i32, _i32, _p,
                                              _i32, _i32, _p, _p, _f32, _p, _p, _i64, _i64, _i64, _p, _i32, _i32, _i32, _i32, _i32,
                                              _i32, _p, _p, _i32, _p]),
    "b200attn_kv_gather": (C.c_int, [_p, _p, _i32, _i64, _i64, _i64, _i64, _p, _i32, _i32, _i32, _i32, _p, _p]),
    "b200attn_plan_bytes": (_sz, [_i32]),
    "b200attn_default_ctas": (C.c_int, [_i32]),
    "b200attn_decode_plan": (C.c_int, [_p, _i32, _i32, _i32, _i32, _p, _i64, _p, _i32, _p, _sz, _p]),
    "b200attn_workspace_bytes": (_sz, [_i32, _i32]),
    "b200attn_decode": (C.c_int, [_p, _i64, _i64, _p, _i64, _i64, _p, _p, _i64, _i64, _i64, _i64, _p, _i64, _p, _i32,
                                  _i32, _i32, _i32, _i32, _i32, _i32, _f32, _p, _p, _i32, _p, _i32, _p, _sz, _p, _p, _i32, _i32, _p]),
    "b200attn_decode_static": (C.c_int, [_p, _i64, _i64, _p, _i64, _i64, _p, _p, _i64, _i64, _i64, _i64, _p, _i64, _p, _i32,
                                         _i32, _i32, _i32, _i32, _i32, _i32, _f32, _p, _p, _i32, _p, _i32, _p]),
    "b200attn_decode_kv_write": (C.c_int, [_p, _i64, _i64, _p, _i64, _i64, _p, _p, _i64, _i64, _i64, _i64, _p, _p, _p, _i64, _i64, _i64,
                                           _i64, _p, _i64, _p, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _f32, _p, _p, _i32, _p, _i32,
                                           _p, _sz, _p, _p, _i32, _i32, _p]),
    "b200attn_decode_per_token_head": (C.c_int, [_p, _i64, _i64, _p, _i64, _i64, _p, _p, _i64, _i64, _i64, _i64, _p, _p, _i64, _i64, _i64,
                                                 _p, _i64, _p, _i32, _i32, _i32, _i32, _i32, _i32, _f32, _p, _i32, _p, _sz, _p, _p,
                                                 _i32, _i32, _p]),
    "b200attn_merge": (C.c_int, [_p, _i32, _p, _p, _i64, _i64, _i32, _i32, _i32, _i32, _i32, _p, _i32, _p, _p]),
    "b200attn_merge_lse": (C.c_int, [_p, _p, _i32, _i64, _i64, _p, _p, _i32, _i32, _i32, _i32, _p]),
    "b200attn_merge_lse_peers": (C.c_int, [_p, _p, _i32, _i64, _i64, _p, _p, _i32, _i32, _i32, _i32, _p]),
    "b200attn_peer_ll_bytes": (_sz, [_i32, _i32, _i32, _i32, _i32]),
    "b200attn_peer_gather_heads": (C.c_int, [_p, _i64, _i64, _i32, _i32, _i32, _i32, _i32, _p, _p, _p, _i32, _i32, _p]),
    "b200attn_peer_combine": (C.c_int, [_p, _p, _i32, _i32, _i32, _i32, _i32, _p, _p, _i32, _i32, _p, _i64, _i64, _p]),
    "b200attn_prefill_tile_tokens": (C.c_int, [_i32, _i32]),
    "b200attn_prefill_plan": (C.c_int, [_p, _i32, _i32, _i32, _i32, _p, _p]),
    "b200attn_prefill": (C.c_int, [_p, _i64, _i64, _p, _i64, _i64, _p, _p, _i64, _i64, _i64, _i64, _p, _i64, _p, _p, _i32, _i32,
                                   _i32, _i32, _i32, _i32, _i32, _i32, _i32, _f32, _p, _p, _i32, _p, _p]),
    "b200attn_set_option": (C.c_int, [C.c_char_p, _i32]),
    "b200attn_debug_set_timeline_buffer": (None, [_p]),
}


def load_library():
    """dlopen the in-tree library once and attach prototypes.  Raises if it has not been built."""
    global _LIB
    if _LIB is not None:
        return _LIB
    with _LOCK:
        if _LIB is not None:
            return _LIB
        path = lib_path()
        if not os.path.exists(path):
            raise B200AttnError(
                f"{path} not found: build it with `python -m ht_vllm_omni_b200._build` "
                "(there is no CPU / PyTorch fallback for this path)"
            )
        lib = C.CDLL(path)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(lib, name)  # AttributeError here = header / library drift
            fn.restype = res
            fn.argtypes = args
        if lib.b200attn_abi_version() != ABI_VERSION:
            raise B200AttnError("libb200attn ABI version mismatch")
        # runtime knobs from the environment: B200ATTN_OPT_<name>=<int> -> b200attn_set_option (see include/b200attn.h)
        for k, v in os.environ.items():
            if k.startswith("B200ATTN_OPT_"):
                if lib.b200attn_set_option(k[len("B200ATTN_OPT_"):].lower().encode(), int(v)) < 0 and int(v) >= 0:
                    pass   # unknown names are reported by b200attn_last_error(); previous value 0 also returns 0
        _LIB = lib
        return lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load_library().b200attn_last_error().decode("utf-8", "replace")
        raise B200AttnError(f"{what} failed ({rc}): {msg}")
