"""-m gpu: the plugin under the real vLLM engine (tests/tools/engine_check.py).  A random-init Qwen3 model of talker shape
(16:2 heads, head_dim 128, 4 layers) is served by `vllm.LLM` -- EngineCore in its own process, the backend found through
the `vllm.general_plugins` entry point, FP8 KV cache, CUDA graphs on -- once with attention_backend=CUSTOM and once with each
of vLLM's own backends.  CUSTOM must agree with TRITON_ATTN on logprobs as closely as vLLM's FLASHINFER backend does."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOOL = os.path.join(ROOT, "tests", "tools", "engine_check.py")


def _arm(arm, work, out, extra=(), timeout=600):
    env = dict(os.environ)
    env.pop("B200ATTN_SEPARATE_KV_UPDATE", None)
    r = subprocess.run([sys.executable, TOOL, "--arm", arm, "--workdir", work, "--out", out, *extra], capture_output=True, text=True,
                       timeout=timeout, env=env, cwd=ROOT)
    return r


def _skip_unless_vllm_starts(r, arm):
    if r.returncode != 0:
        tail = (r.stdout + r.stderr)[-1500:]
        if arm != "custom":
            pytest.skip(f"vLLM's own {arm} engine did not start on this box: {tail[-400:]}")
        raise AssertionError(f"engine with attention_backend=CUSTOM failed:\n{tail}")


@pytest.mark.skipif(os.environ.get("B200ATTN_SKIP_ENGINE_TEST") == "1", reason="B200ATTN_SKIP_ENGINE_TEST=1")
def test_custom_backend_decodes_under_vllm_engine_like_vllms_own_backends(tmp_path):
    assert torch.cuda.is_available()
    pytest.importorskip("vllm")
    work = str(tmp_path)
    outs = {a: os.path.join(work, f"{a}.json") for a in ("custom", "triton", "flashinfer")}
    r = _arm("triton", work, outs["triton"])
    _skip_unless_vllm_starts(r, "triton")
    r = _arm("custom", work, outs["custom"])
    _skip_unless_vllm_starts(r, "custom")
    got = json.load(open(outs["custom"]))
    assert len(got["requests"]) == 8 and all(len(x) == 64 for x in got["requests"])
    r = _arm("flashinfer", work, outs["flashinfer"])
    if r.returncode == 0:
        c = subprocess.run([sys.executable, TOOL, "--compare3", outs["custom"], outs["triton"], outs["flashinfer"]],
                           capture_output=True, text=True, cwd=ROOT)
    else:   # no second native backend on this box: fixed tolerance (fp8 KV: the native backends differ by ~0.3 between themselves)
        c = subprocess.run([sys.executable, TOOL, "--compare", outs["custom"], outs["triton"], "--tol", "0.5"],
                           capture_output=True, text=True, cwd=ROOT)
    print(c.stdout)
    assert c.returncode == 0, c.stdout + c.stderr


@pytest.mark.skipif(os.environ.get("B200ATTN_SKIP_ENGINE_TEST") == "1", reason="B200ATTN_SKIP_ENGINE_TEST=1")
def test_custom_backend_tensor_parallel_2_under_vllm_engine(tmp_path):
    """KV heads sharded over two workers by vLLM itself (tensor_parallel_size=2: one KV head of the talker per GPU)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    pytest.importorskip("vllm")
    work = str(tmp_path)
    a, b = os.path.join(work, "custom_tp2.json"), os.path.join(work, "triton_tp2.json")
    r = _arm("triton", work, b, ("--tp", "2"), timeout=900)
    _skip_unless_vllm_starts(r, "triton")
    r = _arm("custom", work, a, ("--tp", "2"), timeout=900)
    _skip_unless_vllm_starts(r, "custom")
    c = subprocess.run([sys.executable, TOOL, "--compare", a, b, "--tol", "0.5"], capture_output=True, text=True, cwd=ROOT)
    print(c.stdout)
    assert c.returncode == 0, c.stdout + c.stderr
