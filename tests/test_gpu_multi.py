"""-m gpu, needs >= 2 GPUs (skipped otherwise): the multi-GPU data path on real devices, one process per GPU under
torch.distributed.run (NCCL).  Covers what tests/test_parallel_gloo.py can only emulate on the CPU:

  * the sequence-split exchange of a replica group (tp > Hkv) over NCCL (two all-gathers + merge_lse) and over NVLink peer
    memory (parallel.PeerExchange: decode writes (o, lse) into a symmetric-memory buffer, device-side barrier,
    b200attn_merge_lse_peers reads the peers' partials) -- both against the SAME kernel run unsharded on the full context,
    including a sequence shorter than the group (a rank that holds none of its tokens must contribute (0, -inf));
  * the strong-scaled thinker block bench.py reports for --gpus N (tests/tools/sharded_thinker.py) at world size 2.
"""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _torchrun(script, nproc, port, extra=(), timeout=600):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), script, *extra]
    return subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=ROOT)


def _need(n):
    if not torch.cuda.is_available() or torch.cuda.device_count() < n:
        pytest.skip(f"needs {n} GPUs, this box has {torch.cuda.device_count() if torch.cuda.is_available() else 0}")


def test_sequence_split_exchange_matches_unsharded_on_two_gpus():
    _need(2)
    r = _torchrun(os.path.join("tests", "tools", "dcp_check.py"), 2, 29611)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "over NCCL: max |sequence-split - unsharded|" in r.stdout and "over NVLink peer memory" in r.stdout and "over NVLink pushes" in r.stdout, r.stdout[-2000:]


def test_strong_scaled_thinker_block_on_two_gpus():
    _need(2)
    r = _torchrun("bench.py", 2, 29613, extra=("--gpus", "2", "--steps", "3", "--warmup", "3", "--workload", "c1", "--no-cpu-baseline"))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    import json

    line = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    sh = line["sharded"]
    assert "error" not in sh, sh
    assert sh["n_gpus"] == 2 and sh["max_abs_vs_unsharded"] <= sh["tolerance"], sh
    assert 0.3 < sh["efficiency"] < 1.2, sh
