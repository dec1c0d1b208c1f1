"""The plugin under the REAL vLLM engine: a random-init Qwen3 model of talker shape is decoded by
`vllm.LLM` with `attention_backend=CUSTOM` (this repo, discovered through the `vllm.general_plugins` entry point of a
dist-info directory this script writes -- the same discovery a pip install gives) and by the same engine with vLLM's own
TRITON_ATTN backend, both with an FP8 KV cache, CUDA graphs on.  Everything vLLM asks of a backend on the way is exercised
where it bites: validate_configuration, get_kv_cache_shape / stride order inside GPUModelRunner's allocation,
MetadataBuilder.__init__ with the real KVCacheSpec / VllmConfig, get_cudagraph_support, build() per step, forward() inside
captured graphs.

    python tests/tools/engine_check.py --arm custom|triton --out result.json      (one engine per process)
    python tests/tools/engine_check.py --compare a.json b.json

Comparison: greedy decoding of a random-init model is chaotic (near-ties flip), so token equality alone is not the test.
For every request the two runs are compared position by position WHILE their histories agree: the sampled token's logprob
and the top-k logprobs must agree within --tol, and the histories may only part at a position where the top-2 margin is
itself below --tol (a legitimate tie-break difference).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

TALKER = dict(  # Qwen3-TTS talker class defaults (vllm_omni/model_executor/models/qwen3_tts/configuration_qwen3_tts.py:376-398) + head_dim 128
    architectures=["Qwen3ForCausalLM"], model_type="qwen3", hidden_size=1024, intermediate_size=3072, num_hidden_layers=4,
    num_attention_heads=16, num_key_value_heads=2, head_dim=128, vocab_size=3072, max_position_embeddings=4096,
    rms_norm_eps=1e-6, rope_theta=1000000.0, tie_word_embeddings=False, hidden_act="silu", attention_bias=False,
    torch_dtype="bfloat16", initializer_range=0.05, bos_token_id=1, eos_token_id=2)


def write_model(path: str, head_dim: int) -> None:
    """Random-init weights with transformers' own initialiser (seeded), saved as safetensors: both arms load the same file."""
    import torch
    from transformers import Qwen3Config, Qwen3ForCausalLM

    cfg = dict(TALKER)
    cfg["head_dim"] = head_dim
    torch.manual_seed(1234)
    config = Qwen3Config(**{k: v for k, v in cfg.items() if k not in ("architectures", "model_type", "torch_dtype")})
    model = Qwen3ForCausalLM(config).to(torch.bfloat16)
    model.save_pretrained(path, safe_serialization=True)
    with open(os.path.join(path, "config.json")) as f:
        c = json.load(f)
    c["architectures"] = ["Qwen3ForCausalLM"]
    with open(os.path.join(path, "config.json"), "w") as f:
        json.dump(c, f)


def write_dist_info(site: str) -> None:
    """What `pip install .` would leave behind for importlib.metadata: the entry points of pyproject.toml."""
    d = os.path.join(site, "ht_vllm_omni_b200-0.1.0.dist-info")
    os.makedirs(d, exist_ok=True)
    with open(os.path.join(d, "METADATA"), "w") as f:
        f.write("Metadata-Version: 2.1\nName: ht-vllm-omni-b200\nVersion: 0.1.0\n")
    with open(os.path.join(d, "entry_points.txt"), "w") as f:
        f.write("[vllm.general_plugins]\nb200_paged_attention = ht_vllm_omni_b200.plugin:register\n\n"
                "[vllm_omni.general_plugins]\nb200_paged_attention = ht_vllm_omni_b200.plugin:register\n")
    with open(os.path.join(d, "RECORD"), "w") as f:
        f.write("")


def prompts(n: int, vocab: int):
    import numpy as np

    rng = np.random.default_rng(7)
    lens = [5, 37, 130, 260, 17, 64, 300, 511][:n]
    return [rng.integers(3, vocab, size=L).tolist() for L in lens]


def run_arm(args) -> None:
    work = args.workdir or tempfile.mkdtemp(prefix="b200_engine_")
    model_dir = os.path.join(work, f"model_d{args.head_dim}")
    if not os.path.exists(os.path.join(model_dir, "config.json")):
        os.makedirs(model_dir, exist_ok=True)
        write_model(model_dir, args.head_dim)
    site = os.path.join(work, "site")
    write_dist_info(site)
    # children (EngineCore, workers) find the package and its entry point through PYTHONPATH
    os.environ["PYTHONPATH"] = os.pathsep.join([site, ROOT, os.environ.get("PYTHONPATH", "")])
    sys.path[:0] = [site, ROOT]
    os.environ.setdefault("HF_HUB_OFFLINE", "1")
    os.environ.setdefault("TRANSFORMERS_OFFLINE", "1")
    os.environ.setdefault("VLLM_NO_USAGE_STATS", "1")
    os.environ.setdefault("DO_NOT_TRACK", "1")
    os.environ.setdefault("VLLM_LOGGING_LEVEL", "WARNING")
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if args.in_process:
        os.environ["VLLM_ENABLE_V1_MULTIPROCESSING"] = "0"

    from vllm import LLM, SamplingParams

    backend = {"custom": "CUSTOM", "triton": "TRITON_ATTN", "flashinfer": "FLASHINFER"}[args.arm]
    kw = dict(model=model_dir, skip_tokenizer_init=True, dtype="bfloat16", kv_cache_dtype=args.kv_cache_dtype, block_size=16,
              attention_backend=backend, enforce_eager=bool(args.eager), max_model_len=2048, max_num_seqs=8,
              gpu_memory_utilization=0.25, seed=0, tensor_parallel_size=args.tp, enable_prefix_caching=False)
    llm = LLM(**kw)
    sp = SamplingParams(temperature=0.0, max_tokens=args.max_tokens, logprobs=args.topk, detokenize=False, ignore_eos=True)
    ps = prompts(8, TALKER["vocab_size"])
    outs = llm.generate([{"prompt_token_ids": p} for p in ps], sp)
    res = []
    for o in outs:
        c = o.outputs[0]
        steps = []
        for tok, lp in zip(c.token_ids, c.logprobs):
            top = sorted(((int(t), float(v.logprob)) for t, v in lp.items()), key=lambda x: -x[1])
            steps.append({"token": int(tok), "logprob": float(lp[tok].logprob), "top": top[: args.topk]})
        res.append(steps)
    loaded = sorted({l.split()[-1] for l in open("/proc/self/maps") if "libb200attn" in l})
    info = {"arm": args.arm, "backend": backend, "requests": res, "libb200attn_in_api_process": loaded}
    try:   # which backend class the engine really picked (in-process engines only; with a core subprocess it lives there)
        from vllm.v1.attention.backends.registry import AttentionBackendEnum

        info["custom_path"] = AttentionBackendEnum.CUSTOM.get_path() if args.arm == "custom" else None
    except Exception as e:  # noqa: BLE001
        info["custom_path"] = f"{type(e).__name__}: {e}"
    with open(args.out, "w") as f:
        json.dump(info, f)
    print(f"{args.arm}: {len(res)} requests x {args.max_tokens} tokens written to {args.out}")


def compare3(custom: str, triton: str, flashinfer: str, slack: float = 1.5, floor: float = 0.05) -> int:
    """The yardstick for "within the reference's own numerics": the largest logprob gap between vLLM's two native backends
    (FLASHINFER vs TRITON_ATTN) on this model.  CUSTOM must stay within `slack` x that gap (at least `floor`) of TRITON_ATTN,
    and may part from its token history only at near-ties of that size."""
    import contextlib
    import io

    def gap(x, y, tol):
        buf = io.StringIO()
        with contextlib.redirect_stdout(buf):
            rc = compare(x, y, tol)
        return rc, json.loads(buf.getvalue())

    _, ref = gap(flashinfer, triton, 1e9)
    tol = max(floor, slack * ref["max_abs_logprob_diff"])
    rc, ours = gap(custom, triton, tol)
    print(json.dumps({"reference_backends_gap": ref["max_abs_logprob_diff"], "tol": tol, "custom_vs_triton": ours}))
    return rc


def compare(a_path: str, b_path: str, tol: float) -> int:
    a, b = json.load(open(a_path)), json.load(open(b_path))
    worst, compared, agree_tokens, total = 0.0, 0, 0, 0
    bad = []
    for r, (ra, rb) in enumerate(zip(a["requests"], b["requests"])):
        same_history = True
        for i, (sa, sb) in enumerate(zip(ra, rb)):
            total += 1
            if not same_history:
                break
            ta, tb = dict(sa["top"]), dict(sb["top"])
            common = set(ta) & set(tb)
            for t in common:
                worst = max(worst, abs(ta[t] - tb[t]))
                if abs(ta[t] - tb[t]) > tol:
                    bad.append((r, i, t, ta[t], tb[t]))
            compared += 1
            if sa["token"] == sb["token"]:
                agree_tokens += 1
            else:
                same_history = False
                # a parting of ways is legitimate only at a near-tie: each run's choice must be within tol of the other's best
                # (a token outside the other run's top-k has no logprob there, only the bound "below the k-th best": its
                # margin counts when even that bound is a clear one, else the side that does list both tokens decides)
                def margin(s_own, t_own, other_tok):
                    if other_tok in t_own:
                        return s_own["logprob"] - t_own[other_tok], True
                    return s_own["logprob"] - min(t_own.values()), False      # a lower bound on the margin
                ma, known_a = margin(sa, ta, sb["token"]); mb, known_b = margin(sb, tb, sa["token"])
                clear = [m for m in (ma, mb) if m > 2 * tol]
                if clear or not (known_a or known_b):
                    bad.append((r, i, "diverged at a clear margin" if clear else "diverged outside both top-k lists", ma, mb))
    print(json.dumps({"positions_compared_with_identical_history": compared, "tokens_equal": agree_tokens,
                      "positions_total": total, "max_abs_logprob_diff": worst, "tol": tol, "violations": bad[:10]}))
    return 1 if bad else 0


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--arm", choices=["custom", "triton", "flashinfer"])
    ap.add_argument("--out")
    ap.add_argument("--workdir")
    ap.add_argument("--compare", nargs=2)
    ap.add_argument("--compare3", nargs=3, metavar=("CUSTOM", "TRITON", "FLASHINFER"))
    ap.add_argument("--tol", type=float, default=0.05)
    ap.add_argument("--max-tokens", type=int, default=64)
    ap.add_argument("--topk", type=int, default=20)
    ap.add_argument("--head-dim", type=int, default=128)
    ap.add_argument("--tp", type=int, default=1)
    ap.add_argument("--kv-cache-dtype", default="fp8")
    ap.add_argument("--eager", type=int, default=0)
    ap.add_argument("--in-process", type=int, default=0)
    a = ap.parse_args()
    if a.compare3:
        sys.exit(compare3(*a.compare3))
    if a.compare:
        sys.exit(compare(a.compare[0], a.compare[1], a.tol))
    run_arm(a)
