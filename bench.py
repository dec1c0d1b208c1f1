#!/usr/bin/env python
"""bench.py — paged decode-attention HBM GB/s and decode tok/s of the AR-stage hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|thinker|...]

Workload (BASELINE.json configs[1], made concrete in SURVEY.md §8d "C2"): Qwen3-TTS talker streaming decode, FP8-e4m3
KV cache, batch 64, 4k context, (Hq, Hkv, D) = (16, 2, 128), block size 16, 20 layers, synthetic data.
One STEP = one decode step of the attention path of the whole model: the split-KV plan once, then for each of the
20 layers the KV write of the new token (+FP8 quantisation), the split-KV paged decode kernel and the merge kernel.
All 20 per-layer caches (2.7 GB) are distinct, so every step streams far more than the 126 MB L2 (no flush needed).

JSON keys (one line, rank 0): see the task contract.  `value` = algorithmic bytes of a step / device time with inputs
resident in HBM (CUDA-graph replay); `e2e` = the same through the public ops with HOST inputs/outputs (pinned
H2D of q/k/v/metadata and D2H of the attention output inside the timed region); `roofline` = the decode kernel alone,
timed per launch with CUDA events; `cpu_baseline` = the C/OpenMP oracle port on the host cores (bounded sample).
N > 1: KV heads are sharded across ranks (vLLM TP layout): every rank owns Hkv KV heads (and their G*Hkv query heads)
of a model with N*Hkv KV heads, its own cache shard, replicated block tables — weak scaling, no data-path collective
(vLLM's row-parallel o_proj consumes head-sharded outputs; the optional output all-gather is `--allgather`).
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

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (B, L, Hq, Hkv, D, block_size, kv, layers, description)
    "c2": dict(B=64, L=4096, Hq=16, Hkv=2, D=128, bs=16, kv="fp8", layers=20,
               desc="Qwen3-TTS talker streaming decode, FP8 KV, batch 64, 4k context (BASELINE configs[1])"),
    "c2-bf16": dict(B=64, L=4096, Hq=16, Hkv=2, D=128, bs=16, kv="bf16", layers=20, desc="C2 with bf16 KV"),
    "thinker": dict(B=32, L=32768, Hq=32, Hkv=4, D=128, bs=16, kv="fp8", layers=2,
                    desc="Qwen3-Omni thinker decode, FP8 KV, 32k context (BASELINE configs[3], 2 layers resident)"),
    "c1": dict(B=1, L=2048, Hq=16, Hkv=2, D=128, bs=16, kv="bf16", layers=20, desc="talker B1 2k bf16 (configs[0])"),
}


def algorithmic_bytes(lens, Hq, Hkv, D, bs, s_kv):
    """BASELINE.md 'Algorithmic bytes': (decode bytes, kv-write bytes) per layer call."""
    lens = np.asarray(lens, dtype=np.int64)
    B = len(lens)
    dec = int((2 * lens * Hkv * D * s_kv).sum() + 2 * B * Hq * D * 2 + (-(-lens // bs) * 4).sum() + 4 * B)
    wr = int(B * Hkv * D * (2 * 2 + 2 * s_kv) + 8 * B)
    return dec, wr


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:  # noqa: BLE001
        return 6650.0, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------------------------
# clocks sampler (nvidia-smi during the timed region)
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    """Polls NVML (pynvml) every 2 ms from a thread while `active`; falls back to `nvidia-smi -lms 20` when NVML is missing."""
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.samples = []   # (sm_mhz, sm_max_mhz, reasons bitmask)
        self.active = False
        self._stop = False
        self._thread = None
        self._max = None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            # honour CUDA_VISIBLE_DEVICES: map the CUDA ordinal to the NVML device through its UUID
            import torch
            uuid = str(torch.cuda.get_device_properties(self.gpu).uuid)
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
def cpu_oracle_layer(w, reps_budget_s=15.0, max_reps=5, seed=0):
    """Time one layer (KV write + paged decode) of workload `w` on the CPU.  Returns (seconds per layer, cores, sample)."""
    path = os.path.join(ROOT, "oracle", "liboracle.so")
    if not os.path.exists(path):
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle")], check=True, capture_output=True)
    lib = C.CDLL(path)
    try:
        lib.oracle_set_threads(len(os.sched_getaffinity(0)))
    except Exception:  # noqa: BLE001
        pass
    cores = int(lib.oracle_num_threads())
    B, L, Hq, Hkv, D, bs = w["B"], w["L"], w["Hq"], w["Hkv"], w["D"], w["bs"]
    fp8 = w["kv"] == "fp8"
    rng = np.random.default_rng(seed)
    nblk_seq = -(-L // bs)
    nb = B * nblk_seq + 1
    bt = rng.permutation(nb - 1)[: B * nblk_seq].reshape(B, nblk_seq).astype(np.int32)
    lens = np.full(B, L, dtype=np.int32)
    if fp8:   # NaN-free random e4m3 bytes (|x| <= 240)
        kc = (rng.integers(0, 0x78, size=(nb, bs, Hkv, D), dtype=np.uint8) | (rng.integers(0, 2, size=(nb, bs, Hkv, D), dtype=np.uint8) << 7))
        vc = (rng.integers(0, 0x78, size=(nb, bs, Hkv, D), dtype=np.uint8) | (rng.integers(0, 2, size=(nb, bs, Hkv, D), dtype=np.uint8) << 7))
    else:
        kc = (rng.standard_normal((nb, bs, Hkv, D)).astype(np.float32).view(np.uint32) >> 16).astype(np.uint16)
        vc = (rng.standard_normal((nb, bs, Hkv, D)).astype(np.float32).view(np.uint32) >> 16).astype(np.uint16)
    bf = lambda a: (a.astype(np.float32).view(np.uint32) >> 16).astype(np.uint16)  # noqa: E731
    q = bf(rng.standard_normal((B, Hq, D)))
    kn, vn = bf(rng.standard_normal((B, Hkv, D))), bf(rng.standard_normal((B, Hkv, D)))
    slots = (bt[:, (L - 1) // bs].astype(np.int64) * bs + (L - 1) % bs)
    ks = np.full(Hkv, 0.01, np.float32); vs = np.full(Hkv, 0.01, np.float32)
    out = np.zeros((B, Hq, D), np.float32)
    p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731

    def once():
        lib.oracle_kv_write(p(kn), p(vn), p(kc), p(vc), p(slots), B, Hkv, D, bs, 0, int(fp8), p(ks), p(vs), 1)
        lib.oracle_paged_decode(p(q), p(kc), p(vc), p(bt), C.c_int64(bt.shape[1]), p(lens), B, Hq, Hkv, D, bs, 0, int(fp8),
                                C.c_float(D ** -0.5), p(ks), p(vs), 1, p(out))

    once()  # warm-up (page-in)
    times = []
    t_start = time.perf_counter()
    while len(times) < max_reps and (time.perf_counter() - t_start) < reps_budget_s:
        t0 = time.perf_counter(); once(); times.append(time.perf_counter() - t0)
    sample = f"1 of {w['layers']} layers (KV write + paged decode, B={B}, L={L}), {len(times)} reps after 1 warm-up, median"
    return times, cores, sample


def run_reference(args, w):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    s_kv = 1 if w["kv"] == "fp8" else 2
    dec_b, wr_b = algorithmic_bytes([w["L"]] * w["B"], w["Hq"], w["Hkv"], w["D"], w["bs"], s_kv)
    # a reference "step" is a bounded sample: ONE layer of the step on the host cores (1 untimed warm-up inside)
    times, cores, _ = cpu_oracle_layer(w, reps_budget_s=120.0, max_reps=max(1, min(args.steps, 5)))
    t = float(np.median(times))
    gbs = (dec_b + wr_b) / t / 1e9
    sample = f"step = 1 of {w['layers']} layers on the host (oracle/oracle.c, OpenMP), {len(times)} steps, median"
    line = {
        "impl": "reference", "metric": "paged_decode_attn_hbm_gbs", "value": gbs, "unit": "GB/s", "n_gpus": args.gpus,
        "steps": len(times), "warmup": 1, "ms_per_step": t * 1e3 * w["layers"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "fp32 (bf16 activations, e4m3 KV)" if w["kv"] == "fp8" else "fp32 (bf16)",
        "data": "synthetic",
        "config": workload_config(args, w, 0),
        "tok_s": w["B"] / (t * w["layers"]),
        "cpu_baseline": {"value": gbs, "unit": "GB/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": gbs, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference CPU backend (vLLM cpu_attn) is not compiled into the CUDA wheel in this image; this arm times the "
                "oracle's C/OpenMP port of the reference algorithm on all host cores; ms_per_step is extrapolated to all layers",
    }
    emit(json.dumps(line))


def workload_config(args, w, num_ctas):
    return {"workload": f"{args.workload}: {w['desc']}", "batch": w["B"], "context": w["L"], "num_q_heads": w["Hq"],
            "num_kv_heads": w["Hkv"], "head_dim": w["D"], "block_size": w["bs"], "kv_cache_dtype": w["kv"], "layers": w["layers"],
            "parallelism": f"kv-head-sharded tp{args.gpus}" if args.gpus > 1 else "single GPU",
            "l2": "per-step working set (all layer caches) >> 126 MB L2; no flush", "decode_ctas": num_ctas,
            "allgather": args.allgather}


# ------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------
def run_ours(args, w):
    import torch
    import torch.distributed as dist

    import ht_vllm_omni_b200.ops as ops
    from ht_vllm_omni_b200.ops import DecodeWorkspace

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = ops.load_library()

    B, L, Hq, Hkv, D, bs, NL = w["B"], w["L"], w["Hq"], w["Hkv"], w["D"], w["bs"], w["layers"]
    fp8 = w["kv"] == "fp8"
    s_kv = 1 if fp8 else 2
    kv_t = torch.uint8 if fp8 else torch.bfloat16
    lens_np = np.full(B, L, dtype=np.int32)
    g = torch.Generator(device=dev).manual_seed(1234 + rank)
    # ---- per-layer paged caches (HND physical), filled THROUGH the write path from N(0,1) tokens ----
    nblk_seq = -(-L // bs)
    nb = int(B * nblk_seq * 1.1) + 1
    perm = torch.randperm(nb, generator=torch.Generator().manual_seed(99))[: B * nblk_seq].to(torch.int32)
    bt = perm.reshape(B, nblk_seq).contiguous()
    pos = torch.arange(L, dtype=torch.int64)
    slots_all = (bt.to(torch.int64)[:, pos // bs] * bs + (pos % bs)[None, :]).reshape(-1).to(dev)
    d_bt, d_lens = bt.to(dev), torch.from_numpy(lens_np).to(dev)
    k_scale = torch.full((Hkv,), 4.5 / 448.0, dtype=torch.float32, device=dev)   # N(0,1) tokens: |x| <= 4.5 sigma
    v_scale = torch.full((Hkv,), 4.5 / 448.0, dtype=torch.float32, device=dev)
    caches = []
    for _ in range(NL):
        phys = torch.zeros((nb, 2, Hkv, bs, D), dtype=kv_t, device=dev)
        kc, vc = phys.permute(0, 1, 3, 2, 4).unbind(1)
        chunk = 1 << 16
        for s0 in range(0, B * L, chunk):
            n = min(chunk, B * L - s0)
            kk = torch.randn(n, Hkv, D, generator=g, device=dev, dtype=torch.float32).to(torch.bfloat16)
            vv = torch.randn(n, Hkv, D, generator=g, device=dev, dtype=torch.float32).to(torch.bfloat16)
            torch.ops.b200attn.kv_write(kk, vv, kc, vc, slots_all[s0:s0 + n], k_scale, v_scale)
        caches.append((kc, vc))
    torch.cuda.synchronize()
    # ---- per-step device inputs/outputs (the graph reads these fixed buffers) ----
    q = torch.randn(NL, B, Hq, D, generator=g, device=dev, dtype=torch.float32).to(torch.bfloat16)
    kn = torch.randn(NL, B, Hkv, D, generator=g, device=dev, dtype=torch.float32).to(torch.bfloat16)
    vn = torch.randn(NL, B, Hkv, D, generator=g, device=dev, dtype=torch.float32).to(torch.bfloat16)
    out = torch.zeros(NL, B, Hq, D, dtype=torch.bfloat16, device=dev)
    step_slots = (bt.to(torch.int64)[:, (L - 1) // bs] * bs + (L - 1) % bs).to(dev)   # the newest token of each sequence
    ws = DecodeWorkspace(dev, B, kv_t, head_dim=D, num_ctas=args.ctas or None)
    gathered = torch.empty(world, NL, B, Hq, D, dtype=torch.bfloat16, device=dev) if (world > 1 and args.allgather != "none") else None
    sm_scale = D ** -0.5

    def step():
        ws.make_plan(d_lens, Hq, Hkv, d_bt, bs)
        for l in range(NL):
            kc, vc = caches[l]
            torch.ops.b200attn.kv_write(kn[l], vn[l], kc, vc, step_slots, k_scale, v_scale)
            # fresh_tokens=1: a decode step's KV write touched only the newest position of each row (what backend.py passes)
            ws.decode(q[l], kc, vc, d_bt, d_lens, out[l], k_scale, v_scale, sm_scale, fresh_tokens=1)
        if gathered is not None:
            dist.all_gather_into_tensor(gathered, out)

    launches_per_step = 1 + 2 * NL   # plan + per layer: kv_write, decode (split-KV partials merged in-kernel)

    # ---- capture one step into a CUDA graph ----
    side = torch.cuda.Stream(device=dev)
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        step(); step()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    use_graph = not args.no_graph
    if use_graph:
        try:
            with torch.cuda.graph(graph):
                step()
        except Exception as e:  # noqa: BLE001
            if rank == 0:
                print(f"# graph capture failed ({type(e).__name__}: {e}); timing eager launches", file=sys.stderr)
            use_graph = False
            torch.cuda.synchronize()
    run_step = graph.replay if use_graph else step

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local)
    sampler.start()

    # ---- (1) device-resident timing ----
    for _ in range(args.warmup):
        run_step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler.active = True
    e0.record()
    for _ in range(args.steps):
        run_step()
    e1.record()
    barrier()
    sampler.active = False
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_per_step = ms / args.steps

    # ---- (2) end-to-end: host inputs -> pinned H2D -> step -> D2H of the attention output ----
    h_q, h_kn, h_vn = q.cpu().pin_memory(), kn.cpu().pin_memory(), vn.cpu().pin_memory()
    h_slots, h_lens, h_bt = step_slots.cpu().pin_memory(), d_lens.cpu().pin_memory(), d_bt.cpu().pin_memory()
    h_out = torch.empty(out.shape, dtype=out.dtype).pin_memory()
    h2d_bytes = sum(t.numel() * t.element_size() for t in (h_q, h_kn, h_vn, h_slots, h_lens, h_bt))
    d2h_bytes = h_out.numel() * h_out.element_size()

    # The host-buffer call is pipelined in layer groups: group g's inputs travel H2D on a copy stream while group g-1
    # computes, and its outputs travel D2H on a second copy stream while group g+1 computes (PCIe is full duplex).  The
    # whole step -- copies, plan, 2 x NL kernels -- is one CUDA graph with three stream branches.
    NG = int(os.environ.get("B200_E2E_GROUPS", "5"))
    NG = NG if (NG > 0 and NL % NG == 0) else 1
    GL = NL // NG
    s_in, s_out = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)

    def e2e_body():
        main = torch.cuda.current_stream()
        s_in.wait_stream(main)
        s_out.wait_stream(main)
        ev_in = []
        with torch.cuda.stream(s_in):
            step_slots.copy_(h_slots, non_blocking=True); d_lens.copy_(h_lens, non_blocking=True); d_bt.copy_(h_bt, non_blocking=True)
            for gi in range(NG):
                sl = slice(gi * GL, (gi + 1) * GL)
                q[sl].copy_(h_q[sl], non_blocking=True); kn[sl].copy_(h_kn[sl], non_blocking=True); vn[sl].copy_(h_vn[sl], non_blocking=True)
                ev = torch.cuda.Event(); ev.record(s_in); ev_in.append(ev)
        for gi in range(NG):
            main.wait_event(ev_in[gi])
            if gi == 0:
                ws.make_plan(d_lens, Hq, Hkv, d_bt, bs)
            for l in range(gi * GL, (gi + 1) * GL):
                kc, vc = caches[l]
                torch.ops.b200attn.kv_write(kn[l], vn[l], kc, vc, step_slots, k_scale, v_scale)
                ws.decode(q[l], kc, vc, d_bt, d_lens, out[l], k_scale, v_scale, sm_scale, fresh_tokens=1)
            ev = torch.cuda.Event(); ev.record(main)
            with torch.cuda.stream(s_out):
                s_out.wait_event(ev)
                sl = slice(gi * GL, (gi + 1) * GL)
                h_out[sl].copy_(out[sl], non_blocking=True)
        main.wait_stream(s_in)
        main.wait_stream(s_out)

    e2e_graph = None
    if use_graph and gathered is None:
        try:
            with torch.cuda.stream(side):
                e2e_body()
            torch.cuda.synchronize()
            e2e_graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(e2e_graph):
                e2e_body()
        except Exception as e:  # noqa: BLE001
            if rank == 0:
                print(f"# e2e graph capture failed ({type(e).__name__}: {e}); using the serial copy/step/copy path", file=sys.stderr)
            e2e_graph = None
            torch.cuda.synchronize()

    def e2e_step():
        if e2e_graph is not None:
            e2e_graph.replay()
        else:
            q.copy_(h_q, non_blocking=True); kn.copy_(h_kn, non_blocking=True); vn.copy_(h_vn, non_blocking=True)
            step_slots.copy_(h_slots, non_blocking=True); d_lens.copy_(h_lens, non_blocking=True); d_bt.copy_(h_bt, non_blocking=True)
            run_step()
            h_out.copy_(out, non_blocking=True)
        torch.cuda.current_stream().synchronize()   # the caller consumes the result every step

    for _ in range(max(3, args.warmup)):
        e2e_step()
    barrier()
    e0.record()
    for _ in range(args.steps):
        e2e_step()
    e1.record()
    barrier()
    ms_e2e = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms_e2e], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_e2e = float(t.item())
    ms_e2e_per_step = ms_e2e / args.steps

    # ---- (3) roofline leg: the decode kernel alone (split-KV + in-kernel merge, exactly the launch the step makes), NL launches
    # over the NL distinct layer caches captured as one CUDA graph with PDL off (no overlap between neighbours), CUDA events
    # around the replays on the launching stream; duration per launch = elapsed / launches ----
    lib.b200attn_set_option(b"pdl", 0)
    ws.make_plan(d_lens, Hq, Hkv, d_bt, bs)
    torch.cuda.synchronize()

    def decode_only():
        for l in range(NL):
            kc, vc = caches[l]
            ws.decode(q[l], kc, vc, d_bt, d_lens, out[l], k_scale, v_scale, sm_scale)

    reps = max(1, min(args.steps, 10))
    kgraph = torch.cuda.CUDAGraph()
    try:
        with torch.cuda.stream(side):
            decode_only()
        torch.cuda.synchronize()
        with torch.cuda.graph(kgraph):
            decode_only()
        run_k = kgraph.replay
    except Exception:  # noqa: BLE001
        torch.cuda.synchronize()
        run_k = decode_only
    for _ in range(2):
        run_k()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        run_k()
    b.record()
    torch.cuda.synchronize()
    lib.b200attn_set_option(b"pdl", 1)
    evs = [None] * (reps * NL)
    kern_ms = a.elapsed_time(b) / (reps * NL)
    sampler.stop()

    dec_b, wr_b = algorithmic_bytes(lens_np, Hq, Hkv, D, bs, s_kv)
    step_bytes = (dec_b + wr_b) * NL
    peak, peak_src = load_peaks()
    gbs = world * step_bytes / (ms_per_step * 1e-3) / 1e9
    gbs_e2e = world * step_bytes / (ms_e2e_per_step * 1e-3) / 1e9
    kern_gbs = dec_b / (kern_ms * 1e-3) / 1e9

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

    line = {
        "metric": "paged_decode_attn_hbm_gbs", "value": gbs, "unit": "GB/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "fp8_e4m3 KV / bf16 Q,O / tcgen05 kind::f8f6f4 (exact hi+lo e4m3 split of Q and P), fp32 accumulate" if (fp8 and Hq // Hkv <= 8) else
                 "fp8_e4m3 KV / bf16 Q,O / f16 HMMA, fp32 accumulate" if fp8 else "bf16 KV,Q,O / bf16 HMMA, fp32 accumulate",
        "data": "synthetic (N(0,1) tokens written through the KV-write path; random block permutation)",
        "config": workload_config(args, w, ws.grid),
        "tok_s": B / (ms_per_step * 1e-3),
        "pct_of_8tbs": gbs / world / 8000.0 * 100.0, "pct_of_measured_hbm": gbs / world / peak * 100.0,
        "cuda_graph": use_graph,
        "e2e": {"value": gbs_e2e, "unit": "GB/s", "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h_bytes,
                "ms_per_step": ms_e2e_per_step, "tok_s": B / (ms_e2e_per_step * 1e-3),
                "pipeline": f"one CUDA graph, {NG} layer groups: H2D / compute / D2H on three stream branches" if e2e_graph is not None else "serial copy, step, copy"},
        "gpu_launches": launches_per_step * args.steps,
        "roofline": {"bound": "hbm", "achieved": kern_gbs, "peak": peak, "unit": "GB/s", "frac": kern_gbs / peak,
                     "traffic": None, "kernel": ("b200attn::decode_tc_kernel" if (fp8 and Hq // Hkv <= 8) else "b200attn::decode_kernel") + " (split-KV paged flash-decoding, in-kernel merge; CUDA graph of one launch per layer cache, PDL off)",
                     "algorithmic_bytes_per_launch": dec_b, "avg_launch_us": kern_ms * 1e3, "launches_timed": len(evs),
                     "peak_source": peak_src},
        "cpu_baseline": cpu,
        "clocks": sampler.summary(),
    }
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            line["roofline"]["traffic"] = json.load(f).get(args.workload)
    except Exception:  # noqa: BLE001
        pass
    emit(json.dumps(line))
    if world > 1:
        sys.stdout.flush(); sys.stderr.flush()
        os._exit(0)


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
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--ctas", type=int, default=0, help="decode grid override (0 = 2 CTAs/SM)")
    ap.add_argument("--allgather", default="none", choices=["none", "step"], help="NCCL all-gather of head-sharded outputs")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    w = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference(args, w)
    else:
        run_ours(args, w)


if __name__ == "__main__":
    main()
