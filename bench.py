#!/usr/bin/env python
"""bench.py — paged decode-attention HBM GB/s and decode tok/s of the AR-stage hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference|flashinfer|triton] [--workload c2|thinker|...]

Workload (BASELINE.json configs[1], made concrete in SURVEY.md §8d "C2"): Qwen3-TTS talker streaming decode, FP8-e4m3
KV cache, batch 64, 4k context, (Hq, Hkv, D) = (16, 2, 128), block size 16, 20 layers, synthetic data.
One STEP = one decode step of the attention path of the whole model, driven through the vLLM plugin classes exactly as
the reference's runner drives them: `MetadataBuilder.build` once (row buffers + the device-side split-KV plan), then for
each of the 20 layers `Impl.do_kv_cache_update` (KV write + FP8 quantisation) and `Impl.forward` (split-KV paged decode
with the merge in-kernel), replayed as one CUDA graph (`ms_per_step_eager_build`: build() issued eagerly before a graph
of the 40 layer launches, which is what vLLM's FULL cudagraph mode does).
All 20 per-layer caches (2.7 GB) are distinct, so every step streams far more than the 126 MB L2 (no flush needed).

JSON keys (one line, rank 0): see the task contract.
  value     algorithmic bytes of a step / device time, inputs resident in HBM, K steps in one timed region (CUDA events);
            `distribution` = median / p10 / p90 over >= 100 individually timed steps.
  e2e       the same plugin calls with HOST inputs: pinned H2D of q/k/v/metadata and D2H of the attention output inside
            the timed region, every step.
  roofline  the decode kernel alone: one launch per layer cache, launched exactly as the step launches it (programmatic
            dependent launch on), CUDA events around the graph replays; `isolated` = the same with PDL off (no overlap
            between neighbouring launches).
  vs_sm100_reference   the reference's own Blackwell kernels on the same inputs, same method (FlashInfer TRT-LLM-gen with
            bf16 and fp8 query, Triton unified_attention, FA4 for bf16 KV, vLLM's reshape_and_cache_flash).
  cpu_baseline   the C/OpenMP oracle port on the host cores (bounded sample).
N > 1: `value` stays the weak-scaling line (every rank owns Hkv KV heads of a model with N*Hkv KV heads: vLLM's TP layout,
no data-path collective); `sharded` adds the north-star config: the Qwen3-Omni thinker (Hq 32 / Hkv 4, B 32, 32k, fp8)
STRONG-scaled over the N GPUs -- by KV head up to 4, KV head x 2-way sequence split through NVLink peer memory at 8.
--impl reference: the oracle's C/OpenMP port on the host cores (the reference's CPU backend is not in this wheel);
--impl flashinfer / triton: the reference's GPU kernels through the same step (comparison arms, never the product).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time
import types

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
_FI = os.path.join(ROOT, "baseline", "_fi")
if os.path.isdir(_FI):   # FlashInfer's JIT cache prebuilt in-tree (baseline/_fi, git-ignored): no 80 s compile on the box
    os.environ.setdefault("FLASHINFER_WORKSPACE_BASE", _FI)

from tests.tools.workloads import WORKLOADS, algorithmic_bytes, seq_lens_of  # noqa: E402


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:  # noqa: BLE001
        return 6650.0, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------------------------
# clocks sampler (NVML during the timed region)
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    """Polls NVML (pynvml) every 2 ms from a thread while `active`; falls back to `nvidia-smi -lms 20` when NVML is missing."""
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.samples = []   # (sm_mhz, sm_max_mhz, reasons)
        self.active = False
        self._stop = False
        self._thread = None
        self._max = None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            import torch
            uuid = str(torch.cuda.get_device_properties(self.gpu).uuid)   # honour CUDA_VISIBLE_DEVICES
            try:
                h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid).encode())
            except Exception:  # noqa: BLE001
                h = pynvml.nvmlDeviceGetHandleByIndex(self.gpu)
            self._max = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
            bits = {"hw_slowdown": pynvml.nvmlClocksThrottleReasonHwSlowdown,
                    "hw_thermal_slowdown": pynvml.nvmlClocksThrottleReasonHwThermalSlowdown,
                    "sw_thermal_slowdown": pynvml.nvmlClocksThrottleReasonSwThermalSlowdown,
                    "sw_power_cap": pynvml.nvmlClocksThrottleReasonSwPowerCap}

            def poll():
                while not self._stop:
                    if self.active:
                        try:
                            mhz = float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
                            r = int(pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h))
                            self.samples.append((mhz, self._max, [n for n, b in bits.items() if r & b]))
                        except Exception:  # noqa: BLE001
                            pass
                    time.sleep(0.002)
        except Exception:  # noqa: BLE001
            fields = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
                      "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
                      "clocks_event_reasons.sw_power_cap")
            try:
                proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={fields}", "--format=csv,noheader,nounits", "-lms", "20",
                                         "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            except Exception:  # noqa: BLE001
                return
            self._proc = proc

            def poll():
                for line in proc.stdout:
                    if self._stop:
                        break
                    if not self.active:
                        continue
                    f = [x.strip() for x in line.split(",")]
                    try:
                        self.samples.append((float(f[1]), float(f[2]),
                                             [n for n, v in zip(self.NAMES, f[4:8]) if v.lower().startswith("active")]))
                    except Exception:  # noqa: BLE001
                        continue
        self._thread = threading.Thread(target=poll, daemon=True)
        self._thread.start()

    def stop(self):
        self._stop = True
        proc = getattr(self, "_proc", None)
        if proc is not None:
            proc.terminate()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm = [s[0] for s in self.samples]
        reasons = sorted({r for s in self.samples for r in s[2]})
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(s[1] for s in self.samples)), "reasons": reasons,
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle's C/OpenMP port on the host cores (cpu_baseline, and --impl reference)
# ------------------------------------------------------------------------------------------------
def cpu_oracle_layer(w, reps_budget_s=15.0, max_reps=5, seed=0, warmup=1):
    """Time one layer (KV write + paged decode) of workload `w` on the CPU.  Returns (times, cores, sample)."""
    path = os.path.join(ROOT, "oracle", "liboracle.so")
    if not os.path.exists(path):
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle")], check=True, capture_output=True)
    lib = C.CDLL(path)
    try:
        lib.oracle_set_threads(len(os.sched_getaffinity(0)))
    except Exception:  # noqa: BLE001
        pass
    cores = int(lib.oracle_num_threads())
    B, Hq, Hkv, D, bs = w["B"], w["Hq"], w["Hkv"], w["D"], w["bs"]
    lens = seq_lens_of(w)
    fp8 = w["kv"] == "fp8"
    rng = np.random.default_rng(seed)
    nblk = -(-lens.astype(np.int64) // bs)
    nb = int(nblk.sum()) + 1
    perm = rng.permutation(nb - 1).astype(np.int32)
    bt = np.zeros((B, int(nblk.max())), np.int32)
    off = 0
    for s in range(B):
        bt[s, : nblk[s]] = perm[off: off + nblk[s]]
        off += int(nblk[s])
    if fp8:   # NaN-free random e4m3 bytes (|x| <= 240)
        kc = (rng.integers(0, 0x78, size=(nb, bs, Hkv, D), dtype=np.uint8) | (rng.integers(0, 2, size=(nb, bs, Hkv, D), dtype=np.uint8) << 7))
        vc = (rng.integers(0, 0x78, size=(nb, bs, Hkv, D), dtype=np.uint8) | (rng.integers(0, 2, size=(nb, bs, Hkv, D), dtype=np.uint8) << 7))
    else:
        kc = (rng.standard_normal((nb, bs, Hkv, D)).astype(np.float32).view(np.uint32) >> 16).astype(np.uint16)
        vc = (rng.standard_normal((nb, bs, Hkv, D)).astype(np.float32).view(np.uint32) >> 16).astype(np.uint16)
    bf = lambda a: (a.astype(np.float32).view(np.uint32) >> 16).astype(np.uint16)  # noqa: E731
    q = bf(rng.standard_normal((B, Hq, D)))
    kn, vn = bf(rng.standard_normal((B, Hkv, D))), bf(rng.standard_normal((B, Hkv, D)))
    last = lens.astype(np.int64) - 1
    slots = bt[np.arange(B), last // bs].astype(np.int64) * bs + last % bs
    ks = np.full(Hkv, 0.01, np.float32); vs = np.full(Hkv, 0.01, np.float32)
    out = np.zeros((B, Hq, D), np.float32)
    p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731

    def once():
        lib.oracle_kv_write(p(kn), p(vn), p(kc), p(vc), p(slots), B, Hkv, D, bs, 0, int(fp8), p(ks), p(vs), 1)
        lib.oracle_paged_decode(p(q), p(kc), p(vc), p(bt), C.c_int64(bt.shape[1]), p(lens), B, Hq, Hkv, D, bs, 0, int(fp8),
                                C.c_float(D ** -0.5), p(ks), p(vs), 1, p(out))

    for _ in range(max(1, warmup)):
        once()  # warm-up (page-in)
    times = []
    t_start = time.perf_counter()
    while len(times) < max_reps and (time.perf_counter() - t_start) < reps_budget_s:
        t0 = time.perf_counter(); once(); times.append(time.perf_counter() - t0)
    sample = (f"1 of {w['layers']} layers (KV write + paged decode, B={B}, ctx={w['L']}), {len(times)} reps after "
              f"{max(1, warmup)} warm-up, median")
    return times, cores, sample


def workload_config(args, w, num_ctas):
    return {"workload": f"{args.workload}: {w['desc']}", "batch": w["B"], "context": w["L"], "num_q_heads": w["Hq"],
            "num_kv_heads": w["Hkv"], "head_dim": w["D"], "block_size": w["bs"], "kv_cache_dtype": w["kv"], "layers": w["layers"],
            "parallelism": f"kv-head-sharded tp{args.gpus}" if args.gpus > 1 else "single GPU",
            "l2": "per-step working set (all layer caches) >> 126 MB L2; no flush", "decode_ctas": num_ctas,
            "allgather": args.allgather}


def run_reference(args, w):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    s_kv = 1 if w["kv"] == "fp8" else 2
    dec_b, wr_b = algorithmic_bytes(seq_lens_of(w), w["Hq"], w["Hkv"], w["D"], w["bs"], s_kv)
    # a reference "step" is a bounded sample of the step: ONE of its layers on the host cores; --steps / --warmup are honoured
    times, cores, _ = cpu_oracle_layer(w, reps_budget_s=240.0, max_reps=max(1, args.steps), warmup=max(1, args.warmup))
    t = float(np.median(times))
    gbs = (dec_b + wr_b) / t / 1e9
    sample = f"step = 1 of {w['layers']} layers on the host (oracle/oracle.c, OpenMP), {len(times)} steps, median"
    line = {
        "impl": "reference", "metric": "paged_decode_attn_hbm_gbs", "value": gbs, "unit": "GB/s", "n_gpus": args.gpus,
        "steps": len(times), "warmup": max(1, args.warmup), "ms_per_step": t * 1e3 * w["layers"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None,
        "dtype": "fp32 (bf16 activations, e4m3 KV)" if w["kv"] == "fp8" else "fp32 (bf16)", "data": "synthetic",
        "config": workload_config(args, w, 0),
        "tok_s": w["B"] / (t * w["layers"]),
        "cpu_baseline": {"value": gbs, "unit": "GB/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": gbs, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference CPU backend (vLLM cpu_attn) is not compiled into the CUDA wheel in this image; this arm times the "
                "oracle's C/OpenMP port of the reference algorithm on all host cores; ms_per_step is extrapolated to all layers",
    }
    emit(json.dumps(line))


# ------------------------------------------------------------------------------------------------
# helpers shared by the GPU arms
# ------------------------------------------------------------------------------------------------
def capture(body, side):
    """Warm `body` up on a side stream and capture it into a CUDA graph; returns (callable, captured?)."""
    import torch

    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        body(); body()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    try:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            body()
        return g.replay, True
    except Exception as e:  # noqa: BLE001
        print(f"# graph capture failed ({type(e).__name__}: {str(e)[:200]}); timing eager launches", file=sys.stderr)
        torch.cuda.synchronize()
        return body, False


def time_region(run, n, barrier):
    import torch

    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(n):
        run()
    e1.record()
    barrier()
    return e0.elapsed_time(e1)


def max_over_ranks(ms, dev, world):
    if world <= 1:
        return ms
    import torch
    import torch.distributed as dist

    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


class PluginStep:
    """The hot path as vLLM drives it: B200PagedAttentionMetadataBuilder.build -> per layer Impl.do_kv_cache_update +
    Impl.forward ($VLLM/model_executor/layers/attention/attention.py:437-529; caller
    vllm_omni/worker/gpu_ar_model_runner.py:228-298).  Config objects are stand-ins with the fields the classes read."""

    def __init__(self, P, dev):
        import torch
        from vllm.v1.attention.backend import CommonAttentionMetadata

        from ht_vllm_omni_b200 import backend as BK

        w = P.w
        B, Hq, Hkv, D, bs = w["B"], w["Hq"], w["Hkv"], w["D"], w["bs"]
        self.P, self.B, self.Hq, self.Hkv, self.D = P, B, Hq, Hkv, D
        be = BK.B200PagedAttentionBackend
        self.fwd_writes = bool(be.forward_includes_kv_cache_update)
        self.impl = be.get_impl_cls()(Hq, D, P.sm_scale, Hkv, None, None, "fp8" if P.fp8 else "auto")
        self.layer = types.SimpleNamespace(_k_scale=P.k_scale, _v_scale=P.v_scale, _k_scale_float=1.0, _v_scale_float=1.0)
        mc = types.SimpleNamespace(get_num_attention_heads=lambda pc: Hq, get_num_kv_heads=lambda pc: Hkv, get_head_size=lambda: D)
        cfg = types.SimpleNamespace(model_config=mc, parallel_config=None, scheduler_config=types.SimpleNamespace(max_num_batched_tokens=B))
        self.builder = be.get_builder_cls()(types.SimpleNamespace(block_size=bs, dtype=P.kv_t), [f"l{i}" for i in range(P.NL)], cfg, dev)
        self.kv_caches = [ph.permute(0, 1, 3, 2, 4) for ph in P.phys]   # logical [nb, 2, bs, Hkv, D] over HND memory (vLLM's view)
        qsl = torch.arange(B + 1, dtype=torch.int32)
        self.cm = CommonAttentionMetadata(query_start_loc=qsl.to(dev), query_start_loc_cpu=qsl, seq_lens=P.lens, num_reqs=B,
                                          num_actual_tokens=B, max_query_len=1, max_seq_len=int(P.lens_np.max()),
                                          block_table_tensor=P.bt, slot_mapping=P.step_slots)
        self.md = self.builder.build(0, self.cm)
        self.use_inputs(P.q, P.kn, P.vn)
        self.o2 = P.out.view(P.NL, B, Hq * D)

    def use_inputs(self, q, kn, vn):
        """[NL, B, H, D] tensors (or views of one packed qkv activation, as vLLM's qkv split hands them over)"""
        NL, B = self.P.NL, self.B
        self.q2 = q.view(NL, B, self.Hq * self.D); self.k2 = kn.view(NL, B, self.Hkv * self.D); self.v2 = vn.view(NL, B, self.Hkv * self.D)

    def build(self):
        self.md = self.builder.build(0, self.cm)   # eager, before every step / replay -- as vLLM does

    def layers(self, lo=0, hi=None):
        for l in range(lo, self.P.NL if hi is None else hi):
            if not self.fwd_writes:   # attention.py:493: a backend whose forward() includes the KV write gets no separate update call
                self.impl.do_kv_cache_update(self.layer, self.k2[l], self.v2[l], self.kv_caches[l], self.cm.slot_mapping)
            self.impl.forward(self.layer, self.q2[l], self.k2[l], self.v2[l], self.kv_caches[l], self.md, self.o2[l])

    @property
    def grid(self):
        return self.builder.workspace.grid


class OpsStep:
    """Fallback when the vLLM plugin classes cannot be imported on the box: the same launches through torch.ops.b200attn.*."""

    def __init__(self, P, dev):
        import ht_vllm_omni_b200.ops as ops

        w = P.w
        self.P = P
        self.ws = ops.DecodeWorkspace(dev, w["B"], P.kv_t, head_dim=w["D"], group_size=w["Hq"] // w["Hkv"])
        self.use_inputs(P.q, P.kn, P.vn)
        self.build()

    def use_inputs(self, q, kn, vn):
        self.q, self.kn, self.vn = q, kn, vn

    def build(self):
        w = self.P.w
        self.ws.make_plan(self.P.lens, w["Hq"], w["Hkv"], self.P.bt, w["bs"])

    def layers(self, lo=0, hi=None):
        import torch

        P = self.P
        for l in range(lo, P.NL if hi is None else hi):
            kc, vc = P.planes[l]
            torch.ops.b200attn.kv_write(self.kn[l], self.vn[l], kc, vc, P.step_slots, P.k_scale, P.v_scale)
            self.ws.decode(self.q[l], kc, vc, P.bt, P.lens, P.out[l], P.k_scale, P.v_scale, P.sm_scale, fresh_tokens=1)

    @property
    def grid(self):
        return self.ws.grid


# ------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------
def run_ours(args, w):
    import torch
    import torch.distributed as dist

    import ht_vllm_omni_b200.ops as ops
    from tests.tools.workloads import Problem

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = ops.load_library()

    B, Hq, Hkv, D, bs, NL = w["B"], w["Hq"], w["Hkv"], w["D"], w["bs"], w["layers"]
    P = Problem(w, dev, seed=1234 + rank, fill="write")   # caches filled THROUGH the write path from N(0,1) tokens
    try:
        S = PluginStep(P, dev)
        api = ("vLLM plugin classes: B200PagedAttentionMetadataBuilder.build + Impl.forward per layer (forward_includes_kv_cache_update: "
               "the decode launch writes the step's tokens itself)" if S.fwd_writes else
               "vLLM plugin classes: B200PagedAttentionMetadataBuilder.build + Impl.do_kv_cache_update + Impl.forward per layer")
    except Exception as e:  # noqa: BLE001
        print(f"# vLLM plugin classes unavailable ({type(e).__name__}: {str(e)[:200]}); driving torch.ops.b200attn directly", file=sys.stderr)
        S = OpsStep(P, dev)
        api = "torch.ops.b200attn.* (vLLM not importable on this box)"
    gathered = torch.empty(world, NL, B, Hq, D, dtype=torch.bfloat16, device=dev) if (world > 1 and args.allgather != "none") else None

    def step_body():
        # build() is capture-safe (two small copies into the builder's fixed buffers + the device-side plan kernel): the whole
        # step is ONE graph, so the device-timed number does not depend on host enqueue jitter (N ranks share the host cores)
        S.build()
        S.layers()
        if gathered is not None:
            dist.all_gather_into_tensor(gathered, P.out)

    side = torch.cuda.Stream(device=dev)
    step, use_graph = (step_body, False) if args.no_graph else capture(step_body, side)
    # vLLM itself runs build() eagerly before each replay of a graph that holds only the layer hooks: timed as an extra key
    hooks_only, _ = (S.layers, False) if (args.no_graph or gathered is not None) else capture(S.layers, side)

    def step_eager_build():
        S.build()
        hooks_only()

    fwd_writes = bool(getattr(S, "fwd_writes", False)) and P.fp8 and Hq // Hkv <= 8
    # plan + per layer: decode with the KV write and the split-KV merge in-kernel, or kv_write + decode
    launches_per_step = 1 + (1 if fwd_writes else 2) * NL

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local)
    sampler.start()

    # ---- (1) device-resident timing: K steps in one region ----
    for _ in range(args.warmup):
        step()
    sampler.active = True
    ms = max_over_ranks(time_region(step, args.steps, barrier), dev, world)
    sampler.active = False
    ms_per_step = ms / args.steps
    ms_eager_build = max_over_ranks(time_region(step_eager_build, args.steps, barrier), dev, world) / args.steps
    # distribution over >= 100 individually timed steps (20 warm-up + 100 timed, median and p10 / p90)
    n_dist = max(100, args.steps)
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n_dist)]
    barrier()
    for a, b in evs:
        a.record(); step(); b.record()
    barrier()
    per = np.array([a.elapsed_time(b) for a, b in evs])
    dist_ms = {"steps": n_dist, "median_ms": float(np.median(per)), "p10_ms": float(np.percentile(per, 10)),
               "p90_ms": float(np.percentile(per, 90)), "note": "each step bracketed by its own CUDA events (serialises host enqueue and device)"}

    # ---- (2) end-to-end: host inputs -> pinned H2D -> build + layer hooks -> D2H of the attention output, every step ----
    # The step's activations sit in host memory the way the model produces them: one packed qkv row per token and layer
    # ([T, (Hq + 2 Hkv) * D], $VLLM/model_executor/models/qwen3.py:153 `qkv.split`), so a layer group is ONE H2D copy and the
    # hooks read q / k / v as strided views of it -- every copy node of a graph costs ~7 us of DMA set-up on this box.
    h_qkv = torch.cat([P.q, P.kn, P.vn], dim=2).cpu().pin_memory()
    d_qkv = torch.empty(h_qkv.shape, dtype=h_qkv.dtype, device=dev)
    S.use_inputs(d_qkv[:, :, :Hq], d_qkv[:, :, Hq:Hq + Hkv], d_qkv[:, :, Hq + Hkv:])
    h_slots, h_lens, h_bt = P.step_slots.cpu().pin_memory(), P.lens.cpu().pin_memory(), P.bt.cpu().pin_memory()
    h_out = torch.empty(P.out.shape, dtype=P.out.dtype).pin_memory()
    h2d_bytes = sum(t.numel() * t.element_size() for t in (h_qkv, h_slots, h_lens, h_bt))
    d2h_bytes = h_out.numel() * h_out.element_size()
    # The layer part is pipelined in layer groups: group g's q/k/v travel H2D on a copy stream while earlier groups compute, and
    # its outputs travel D2H on a second copy stream while later groups compute (PCIe is full duplex); the whole call is one
    # CUDA graph with three stream branches.  The input branch forks right after the metadata copies, beside build().
    # B200_E2E_GROUPS = a group count (equal sizes) or a comma list of group sizes.
    def group_sizes(spec):
        if "," in spec:
            sizes = [int(x) for x in spec.split(",")]
            return sizes if (sum(sizes) == NL and min(sizes) > 0) else [NL]
        if spec:
            n = int(spec)
            return [NL // n] * n if (n > 0 and NL % n == 0) else [NL]
        if NL < 12:
            return [NL]
        # short first and last groups: the first group's H2D and the last group's D2H are the only copies nothing hides
        # (c2, 20 layers: 5 x 4 -> 623 us, [1, 3, 4, 4, 4, 3, 1] -> 614 us, 10 x 2 -> 646 us, one group -> 772 us per step)
        mid = NL - 8
        return [1, 3] + [4] * (mid // 4) + ([mid % 4] if mid % 4 else []) + [3, 1]

    s_in, s_out = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)

    def make_e2e_body(sizes):
        bounds = [0]
        for n in sizes:
            bounds.append(bounds[-1] + n)

        def e2e_body():
            main = torch.cuda.current_stream()
            # step metadata first (build() reads it; the copy engine serves requests in order), then the qkv copies on their
            # own branch beside build() -- capture-safe: small copies + the device-side plan
            P.step_slots.copy_(h_slots, non_blocking=True); P.lens.copy_(h_lens, non_blocking=True); P.bt.copy_(h_bt, non_blocking=True)
            s_in.wait_stream(main)
            ev_in = []
            with torch.cuda.stream(s_in):
                for gi in range(len(sizes)):
                    sl = slice(bounds[gi], bounds[gi + 1])
                    d_qkv[sl].copy_(h_qkv[sl], non_blocking=True)
                    ev = torch.cuda.Event(); ev.record(s_in); ev_in.append(ev)
            S.build()
            s_out.wait_stream(main)
            for gi in range(len(sizes)):
                main.wait_event(ev_in[gi])
                S.layers(bounds[gi], bounds[gi + 1])
                ev = torch.cuda.Event(); ev.record(main)
                with torch.cuda.stream(s_out):
                    s_out.wait_event(ev)
                    sl = slice(bounds[gi], bounds[gi + 1])
                    h_out[sl].copy_(P.out[sl], non_blocking=True)
            main.wait_stream(s_in)
            main.wait_stream(s_out)
        return e2e_body

    def measure_e2e(sizes):
        body = make_e2e_body(sizes)
        replay, graphed = (body, False)
        if use_graph and gathered is None:
            replay, graphed = capture(body, side)

        def e2e_step():
            replay()
            torch.cuda.current_stream().synchronize()   # the caller consumes the result every step

        for _ in range(max(3, args.warmup)):
            e2e_step()
        return max_over_ranks(time_region(e2e_step, args.steps, barrier), dev, world), graphed

    if os.environ.get("B200_E2E_SWEEP"):   # tuning aid: "5;20;1,1,4,4,4,4,1,1;..." -> one stderr line per schedule
        for spec in os.environ["B200_E2E_SWEEP"].split(";"):
            ms_try, _ = measure_e2e(group_sizes(spec))
            if rank == 0:
                print(f"[e2e sweep] groups {group_sizes(spec)}: {ms_try / args.steps * 1e3:.1f} us/step", file=sys.stderr, flush=True)
    e2e_sizes = group_sizes(os.environ.get("B200_E2E_GROUPS", ""))
    NG = len(e2e_sizes)
    ms_e2e, e2e_graphed = measure_e2e(e2e_sizes)
    ms_e2e_per_step = ms_e2e / args.steps
    S.use_inputs(P.q, P.kn, P.vn)

    # ---- (3) roofline leg: the decode kernel alone, NL launches over the NL distinct layer caches, one CUDA graph ----
    ws = ops.DecodeWorkspace(dev, B, P.kv_t, head_dim=D, group_size=Hq // Hkv)
    ws.make_plan(P.lens, Hq, Hkv, P.bt, bs)
    torch.cuda.synchronize()

    def decode_only():
        for l in range(NL):
            kc, vc = P.planes[l]
            if fwd_writes:   # the launch the step issues: decode that also stores the rows' new tokens
                ws.decode_kv_write(P.q[l], P.kn[l], P.vn[l], P.step_slots, kc, vc, P.bt, P.lens, P.out[l], P.k_scale, P.v_scale,
                                   P.sm_scale, fresh_tokens=1)
            else:
                ws.decode(P.q[l], kc, vc, P.bt, P.lens, P.out[l], P.k_scale, P.v_scale, P.sm_scale, fresh_tokens=1)

    reps = max(3, min(args.steps, 20))
    kern = {}
    for pdl in (1, 0):
        lib.b200attn_set_option(b"pdl", pdl)
        run_k, _ = capture(decode_only, side)
        for _ in range(3):
            run_k()
        kern[pdl] = time_region(run_k, reps, lambda: torch.cuda.synchronize()) / (reps * NL)   # ms per launch
    lib.b200attn_set_option(b"pdl", 1)
    sampler.stop()
    ours_out = P.out.clone()

    dec_b, wr_b = P.dec_bytes, P.wr_bytes
    step_bytes = (dec_b + wr_b) * NL
    peak, peak_src = load_peaks()
    gbs = world * step_bytes / (ms_per_step * 1e-3) / 1e9
    gbs_e2e = world * step_bytes / (ms_e2e_per_step * 1e-3) / 1e9
    kern_gbs = dec_b / (kern[1] * 1e-3) / 1e9
    kern_iso_gbs = dec_b / (kern[0] * 1e-3) / 1e9
    fp8 = P.fp8
    tc = fp8 and Hq // Hkv <= 8

    # ---- (4) the reference's own sm_100 kernels on the same inputs (N = 1, rank 0) ----
    sm100 = None
    if world == 1 and not args.no_sm100_ref:
        try:
            from tests.tools import sm100_reference as R

            arms = R.compare_all(P.phys, P.q, P.bt, P.lens, P.sm_scale, P.k_scale, P.v_scale, dec_b, wr_b, kn=P.kn, vn=P.vn,
                                 slots=P.step_slots, ours_out=ours_out)
            sm100 = {"method": "every arm: CUDA graph of one launch per layer cache, CUDA events, each library's default launch attributes "
                               "(FlashInfer enables programmatic dependent launch on sm_100 by itself; so does our step); ratio_vs_ours = arm "
                               "time / our time (`ours_us`, PDL on), > 1 means ours is faster", "ours_us": kern[1] * 1e3,
                     "ours_isolated_pdl_off_us": kern[0] * 1e3}
            # the reference's whole per-layer path next to ours (what a layer of the model pays for this path)
            fi8, qq, wrr = arms.get("flashinfer_trtllm_gen_q_fp8", {}), arms.get("vllm_query_fp8_quant", {}), arms.get("vllm_reshape_and_cache_flash", {})
            if all("us_per_layer" in a for a in (fi8, qq, wrr)):
                sm100["layer_path_us"] = {
                    "reference_flashinfer_fp8_query": fi8["us_per_layer"] + qq["us_per_layer"] + wrr["us_per_layer"],
                    "reference_parts": "trtllm_batch_decode (fp8 q) + scaled_fp8_quant(q) + reshape_and_cache_flash, each timed in its own graph",
                    "ours": ms_per_step * 1e3 / NL,
                    "ours_parts": "ms_per_step / layers (plan included): " + ("decode launch that also writes the KV" if fwd_writes else "kv_write + decode")}
            for k, v in arms.items():
                if "us_per_layer" in v and not k.startswith("vllm_"):
                    v["ratio_vs_ours"] = v["us_per_layer"] / (kern[1] * 1e3)
                sm100[k] = v
        except Exception as e:  # noqa: BLE001
            sm100 = {"error": f"{type(e).__name__}: {str(e)[:300]}"}

    # ---- (5) N > 1: the north-star sharded config, strong-scaled ----
    sharded = None
    if world > 1 and not args.no_sharded:
        try:
            from tests.tools.sharded_thinker import run_sharded

            del P.phys, P.planes
            torch.cuda.empty_cache()
            sharded = run_sharded(dev, rank, world, lib)
        except Exception as e:  # noqa: BLE001
            sharded = {"error": f"{type(e).__name__}: {str(e)[:300]}"}

    if world > 1:
        # every rank is past its last collective: leave without tearing NCCL down (communicators referenced by a captured
        # graph can make destroy_process_group block); rank 0 prints first
        dist.barrier()
        torch.cuda.synchronize()
    if rank != 0:
        sys.stdout.flush(); sys.stderr.flush()
        os._exit(0)

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        times, cores, sample = cpu_oracle_layer(w)
        t_cpu = float(np.median(times))
        cpu = {"value": (dec_b + wr_b) / t_cpu / 1e9, "unit": "GB/s", "cores": cores, "kind": "port", "sample": sample,
               "tok_s": B / (t_cpu * NL)}

    kname = ("b200attn::decode_tc_kernel" if tc else "b200attn::decode_kernel")
    line = {
        "metric": "paged_decode_attn_hbm_gbs", "value": gbs, "unit": "GB/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "fp8_e4m3 KV / bf16 Q,O / tcgen05 kind::f8f6f4 (exact hi+lo e4m3 split of Q and P), fp32 accumulate" if tc else
                 "fp8_e4m3 KV / bf16 Q,O / f16 HMMA, fp32 accumulate" if fp8 else "bf16 KV,Q,O / bf16 HMMA, fp32 accumulate",
        "data": "synthetic (N(0,1) tokens written through the KV-write path; random block permutation)",
        "config": workload_config(args, w, S.grid),
        "api": api,
        "tok_s": B / (ms_per_step * 1e-3),
        "pct_of_8tbs": gbs / world / 8000.0 * 100.0, "pct_of_measured_hbm": gbs / world / peak * 100.0,
        "cuda_graph": use_graph,
        "ms_per_step_eager_build": ms_eager_build,   # build() issued eagerly before each replay of a layers-only graph (vLLM's flow)
        "distribution": dist_ms,
        "e2e": {"value": gbs_e2e, "unit": "GB/s", "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h_bytes,
                "ms_per_step": ms_e2e_per_step, "tok_s": B / (ms_e2e_per_step * 1e-3), "api": api,
                "pipeline": (f"one CUDA graph: metadata H2D, build() beside the first packed-qkv H2D, then {NG} layer groups of {e2e_sizes} layers with H2D / layer hooks / D2H on three stream branches"
                             if e2e_graphed else "metadata H2D, build(), copies and layer hooks issued eagerly")},
        "gpu_launches": launches_per_step * args.steps,
        "roofline": {"bound": "hbm", "achieved": kern_gbs, "peak": peak, "unit": "GB/s", "frac": kern_gbs / peak,
                     "traffic": None,
                     "kernel": kname + (" (split-KV paged flash-decoding, KV write and merge in-kernel)" if fwd_writes else
                                        " (split-KV paged flash-decoding, merge in-kernel)") + "; CUDA graph of one launch per layer cache, launched "
                                       "as the step launches it (programmatic dependent launch on)",
                     "algorithmic_bytes_per_launch": dec_b, "avg_launch_us": kern[1] * 1e3, "launches_timed": reps * NL,
                     "pct_of_8tbs": kern_gbs / 8000.0 * 100.0,
                     "isolated": {"note": "same graph with PDL off: no overlap between neighbouring launches", "avg_launch_us": kern[0] * 1e3,
                                  "achieved": kern_iso_gbs, "frac": kern_iso_gbs / peak, "pct_of_8tbs": kern_iso_gbs / 8000.0 * 100.0},
                     "peak_source": peak_src},
        "vs_sm100_reference": sm100,
        "cpu_baseline": cpu,
        "clocks": sampler.summary(),
    }
    if sharded is not None:
        line["sharded"] = sharded
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            tr = json.load(f).get(args.workload)
            if isinstance(tr, dict):
                line["roofline"]["traffic"] = tr.get("dram_bytes_per_launch")
                line["roofline"]["traffic_source"] = {k: v for k, v in tr.items() if k != "dram_bytes_per_launch"}
            else:
                line["roofline"]["traffic"] = tr
    except Exception:  # noqa: BLE001
        pass
    emit(json.dumps(line))
    if world > 1:
        sys.stdout.flush(); sys.stderr.flush()
        os._exit(0)


# ------------------------------------------------------------------------------------------------
# comparison arms: the reference's GPU kernels through the same step
# ------------------------------------------------------------------------------------------------
def run_sm100(args, w):
    import torch

    from tests.tools import sm100_reference as R
    from tests.tools.workloads import Problem

    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    P = Problem(w, dev, seed=1234, fill="write")
    NL, B = P.NL, w["B"]
    arm = args.impl
    if arm == "flashinfer":
        dec = R.time_flashinfer(P.phys, P.q, P.bt, P.lens, int(P.lens_np.max()), P.sm_scale, P.k_scale[0].item(), P.v_scale[0].item(),
                                P.dec_bytes, q_fp8=P.fp8 and not args.fi_q_bf16)
    else:
        dec = R.time_triton(P.planes, P.q, P.bt, P.lens, int(P.lens_np.max()), P.sm_scale, P.k_scale, P.v_scale, P.dec_bytes, w["Hkv"])
    wr = R.time_vllm_kv_write(P.planes, P.kn, P.vn, P.step_slots, P.k_scale, P.v_scale, P.wr_bytes)
    if "error" in dec or "error" in wr:
        emit(json.dumps({"impl": arm, "unavailable": (dec.get("error") or wr.get("error"))[:300]}))
        return
    dec.pop("out", None)
    us = dec["us_per_layer"] + wr["us_per_layer"]
    ms_per_step = us * NL * 1e-3
    step_bytes = (P.dec_bytes + P.wr_bytes) * NL
    line = {"impl": arm, "metric": "paged_decode_attn_hbm_gbs", "value": step_bytes / (ms_per_step * 1e-3) / 1e9, "unit": "GB/s", "n_gpus": 1,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "data": "synthetic", "config": workload_config(args, w, 0), "tok_s": B / (ms_per_step * 1e-3),
            "decode": dec, "kv_write": wr,
            "note": "comparison arm: the reference's own sm_100 kernels (decode and vLLM's reshape_and_cache_flash timed as separate CUDA "
                    "graphs of one launch per layer cache and summed); never part of the product path"}
    emit(json.dumps(line))


_REAL_STDOUT = None


def emit(text: str) -> None:
    """The ONE JSON line goes to the real stdout; everything else a library prints there (NCCL's version banner) was
    diverted to stderr by main()."""
    if _REAL_STDOUT is None:
        print(text, flush=True)
    else:
        os.write(_REAL_STDOUT, (text + "\n").encode())


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "flashinfer", "triton"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--allgather", default="none", choices=["none", "step"], help="NCCL all-gather of head-sharded outputs")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-sm100-ref", action="store_true", help="skip timing the reference's own Blackwell kernels")
    ap.add_argument("--no-sharded", action="store_true", help="N > 1: skip the strong-scaled thinker block")
    ap.add_argument("--fi-q-bf16", action="store_true", help="--impl flashinfer: keep the query in bf16 (default: fp8 query, as vLLM does)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    w = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference(args, w)
    elif args.impl == "ours":
        run_ours(args, w)
    else:
        run_sm100(args, w)


if __name__ == "__main__":
    main()
