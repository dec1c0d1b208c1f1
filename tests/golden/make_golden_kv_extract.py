"""Golden vectors for the KV hand-off gather from the REFERENCE's own code.

`vllm_omni` does not import in this container (absent dependencies), so the two functions on the path are
taken from the reference's source files with `ast` and executed unmodified in a namespace that provides only torch, a
logger and a stand-in for the KVCacheTransferData dataclass (a plain record of the same fields):
  * normalize_layer_kv                       /root/reference/vllm_omni/distributed/omni_connectors/utils/kv_utils.py:13-85
  * OmniKVTransferManager._extract_kv_cache  /root/reference/vllm_omni/distributed/omni_connectors/kv_transfer_manager.py:224-301
Inputs + outputs go to tests/golden/ref_kv_extract.npz.   Run (CPU, seconds):  python tests/golden/make_golden_kv_extract.py
"""
import ast
import logging
import os
from typing import Any

import numpy as np
import torch

REF = "/root/reference/vllm_omni/distributed/omni_connectors"
HERE = os.path.dirname(os.path.abspath(__file__))


def grab(path, name, cls=None):
    tree = ast.parse(open(path).read())
    body = tree.body
    if cls is not None:
        body = next(n for n in body if isinstance(n, ast.ClassDef) and n.name == cls).body
    fn = next(n for n in body if isinstance(n, ast.FunctionDef) and n.name == name)
    return ast.unparse(fn)


class KVCacheTransferData:   # stand-in: same constructor fields as the reference's dataclass
    def __init__(self, request_id, layer_blocks, block_ids, metadata):
        self.request_id, self.layer_blocks, self.block_ids, self.metadata = request_id, layer_blocks, block_ids, metadata


ns = {"torch": torch, "Any": Any, "logger": logging.getLogger("ref"), "KVCacheTransferData": KVCacheTransferData,
      "LayerKV": Any}
exec(grab(os.path.join(REF, "utils", "kv_utils.py"), "normalize_layer_kv"), ns)
exec(grab(os.path.join(REF, "kv_transfer_manager.py"), "_extract_kv_cache", cls="OmniKVTransferManager"), ns)


def main():
    torch.manual_seed(20261104)
    nb, bs, H, D, L = 12, 16, 2, 64, 2
    out = {}
    cases = {
        "fp8_nb2": dict(dtype=torch.uint8, stacked="nb2", ids=[7, 2, 9, 0], seq_len=50),
        "bf16_2nb": dict(dtype=torch.bfloat16, stacked="2nb", ids=[3, 11, 5], seq_len=48),
        "fp8_invalid_ids": dict(dtype=torch.uint8, stacked="nb2", ids=[4, 99, -1, 6, 1], seq_len=1000),
        "bf16_tuple_short": dict(dtype=torch.bfloat16, stacked="tuple", ids=[10, 1], seq_len=1),
    }
    for name, c in cases.items():
        caches = []
        for _ in range(L):
            if c["dtype"] == torch.uint8:
                t = torch.randint(0, 256, (nb, 2, bs, H, D), dtype=torch.uint8)
            else:
                t = torch.randn(nb, 2, bs, H, D).to(c["dtype"])
            caches.append(t)
        layer_kv = [t if c["stacked"] == "nb2" else (t.transpose(0, 1).contiguous() if c["stacked"] == "2nb" else (t[:, 0], t[:, 1]))
                    for t in caches]
        r = ns["_extract_kv_cache"](None, "req", c["ids"], c["seq_len"], layer_kv, bs, "auto")
        raw = lambda t: t.contiguous().view(torch.uint8).numpy() if t.dtype == torch.uint8 else t.contiguous().view(torch.int16).numpy().view(np.uint16)  # noqa: E731
        out[name + "_caches"] = np.stack([raw(t) for t in caches])
        out[name + "_ids"] = np.array(c["ids"], dtype=np.int64)
        out[name + "_seq_len"] = np.array(c["seq_len"])
        out[name + "_k"] = np.stack([raw(t) for t in r.layer_blocks["key_cache"]])
        out[name + "_v"] = np.stack([raw(t) for t in r.layer_blocks["value_cache"]])
        out[name + "_layout"] = np.array(c["stacked"])
    none_case = ns["_extract_kv_cache"](None, "req", [99, -5], 10, [torch.zeros(nb, 2, bs, H, D)], bs, "auto")
    out["all_invalid_is_none"] = np.array(none_case is None)
    np.savez_compressed(os.path.join(HERE, "ref_kv_extract.npz"), **out)
    print("wrote ref_kv_extract.npz", {k: v.shape for k, v in out.items() if k.endswith("_k")})


if __name__ == "__main__":
    main()
