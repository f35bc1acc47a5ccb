#!/usr/bin/env python
"""Benchmark of the attention forward / split-KV decode hot path (bench contract of the build driver).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c3|c2|c4|c5]

Prints ONE JSON line.  A step is one pass of the hot path over one batch of synthetic input:

    c3 (default, the north-star prefill target): causal forward, bf16, B=8 H=32 N=8192 d=128
    c2: non-causal forward, fp16, B=4 H=16 N=4096 d=64
    c4: split-KV decode, bf16, B=8, 32 q-heads / 8 kv-heads, KV length 32768, d=128 (1 GiB of KV)
    c5: causal forward, bf16, B=1 H=32 N=65536 d=128

``value`` is measured with the inputs resident in HBM (CUDA events on the launching stream, max over
ranks); ``e2e`` is the same metric through the public API with pinned host buffers, host<->device copies
inside the timed region.  Multi-GPU: one process per GPU (torchrun); the (batch x head) units are
independent, so the units of ONE problem are cut over the ranks with no collective in the timed path (strong scaling:
C3 at N=8 is one batch entry per GPU, C5 four heads per GPU).
``--impl reference`` times the reference's own CPU path (the naive softmax(QK^T/sqrt d)V of its Python
example, restated in oracle/naive.py) on the host cores, on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    #        kind      dtype             B  Hq  Hkv  S      D    causal
    "c3": ("fwd", torch.bfloat16, 8, 32, 32, 8192, 128, True),
    "c2": ("fwd", torch.float16, 4, 16, 16, 4096, 64, False),
    "c5": ("fwd", torch.bfloat16, 1, 32, 32, 65536, 128, True),
    "c4": ("decode", torch.bfloat16, 8, 32, 8, 32768, 128, False),
}


def workload_units(name):
    kind, dtype, B, Hq, Hkv, S, D, causal = WORKLOADS[name]
    if kind == "fwd":
        flops = 4.0 * B * Hq * S * S * D * (0.5 if causal else 1.0)       # FlashAttention convention
        return {"kind": kind, "amount": flops / 1e12, "unit": "TFLOP/s", "metric": "attn_fwd_tflops",
                "bound": "tensor"}
    kv_bytes = 2.0 * B * S * Hkv * D * 2                                   # K and V, read once
    return {"kind": kind, "amount": kv_bytes / 1e9, "unit": "GB/s", "metric": "decode_kv_gbs", "bound": "hbm"}


def algorithmic_bytes(name):
    kind, dtype, B, Hq, Hkv, S, D, causal = WORKLOADS[name]
    if kind == "fwd":
        return 2.0 * B * S * D * (2 * Hq + 2 * Hkv) + 4.0 * B * Hq * S      # Q, O, K, V once + fp32 LSE
    return 2.0 * B * S * Hkv * D * 2                                          # K and V once


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return {"hbm": float(p["hbm_gbs"]), "tensor": float(p["bf16_tflops"]),
                "tensor_sustained": float(p.get("bf16_tflops_sustained", p["bf16_tflops"])), "source": "MEASURED_PEAKS.json"}
    return {"hbm": 6650.0, "tensor": 1590.0, "tensor_sustained": 1590.0, "source": "fallback (B200_PROFILING.md)"}


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region.  NVML in a thread (5 ms period) so that
    even a 40 ms region gets samples; `nvidia-smi -lms 100` (B200_PROFILING.md recipe) is the fallback."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []
        self.sm, self.reasons, self.max_mhz, self.power = [], set(), None, []
        self._stop = threading.Event()
        self._thread = None
        self.source = None

    def _nvml_loop(self, nv, h):
        bits = {"hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
                "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
                "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
                "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4)}
        get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
            getattr(nv, "nvmlDeviceGetCurrentClocksThrottleReasons")
        while not self._stop.is_set():
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                self.power.append(nv.nvmlDeviceGetPowerUsage(h) / 1e3)
                r = int(get_reasons(h))
                for name, bit in bits.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:  # noqa: BLE001
                pass
            time.sleep(0.005)

    def start(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = int(vis.split(",")[self.index]) if vis and vis.split(",")[self.index].isdigit() else self.index
            h = nv.nvmlDeviceGetHandleByIndex(idx)
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            self.source = "nvml"
            self._thread = threading.Thread(target=self._nvml_loop, args=(nv, h), daemon=True)
            self._thread.start()
            return
        except Exception:  # noqa: BLE001
            self.source = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.source = "nvidia-smi"
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.source == "nvml":
            self._stop.set()
            self._thread.join(timeout=1.0)
            return {"sm_mhz": statistics.median(self.sm) if self.sm else None, "sm_max_mhz": self.max_mhz,
                    "reasons": sorted(self.reasons), "samples": len(self.sm), "source": "nvml",
                    "power_w_max": max(self.power) if self.power else None}
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "source": "nvidia-smi"}


def make_inputs(shape, device, seed):
    """Synthetic N(0,1) inputs of a (kind, dtype, B, Hq, Hkv, S, D, causal) problem: drawn on the device, mirrored into
    pinned host buffers for the e2e leg."""
    kind, dtype, B, Hq, Hkv, S, D, causal = shape
    g = torch.Generator(device=device).manual_seed(seed)

    def rnd(shp):
        t = torch.empty(shp, dtype=dtype, device=device)
        flat = t.view(-1)
        step = 1 << 26
        for i in range(0, flat.numel(), step):
            n = min(step, flat.numel() - i)
            flat[i:i + n] = torch.randn(n, generator=g, device=device, dtype=torch.float32).to(dtype)
        return t
    if kind == "fwd":
        dev_t = {"q": rnd((B, S, Hq, D)), "k": rnd((B, S, Hkv, D)), "v": rnd((B, S, Hkv, D))}
    else:
        dev_t = {"q": rnd((B, Hq, D)), "k": rnd((B, S, Hkv, D)), "v": rnd((B, S, Hkv, D))}
        dev_t["seqlens"] = torch.full((B,), S, dtype=torch.int32, device=device)
    host = {}
    for k_, v_ in dev_t.items():
        h = torch.empty(v_.shape, dtype=v_.dtype, pin_memory=True)
        h.copy_(v_)
        host[k_] = h
    return host, dev_t


def workload_config(name, world):
    """The `config` object of the JSON line: the same for both arms (it names the workload, not the implementation)."""
    kind, dtype, B, Hq, Hkv, S, D, causal = WORKLOADS[name]
    per = B * Hkv // max(1, world)
    in_mb = algorithmic_bytes(name) / max(1, world) / 1e6
    return {"workload": f"{name}: {kind} B={B} Hq={Hq} Hkv={Hkv} N={S} d={D} causal={causal}",
            "l2": (f"{in_mb:.0f} MB of tensors per rank, larger than the 126 MB L2 (no flush)" if in_mb > 126 * 1.5 else
                   f"{in_mb:.0f} MB of tensors per rank: part of it can stay in the 126 MB L2 between steps (no flush)"),
            "parallelism": f"one problem per step; its {B * Hkv} (batch x kv-head) units cut over {world} rank(s), "
                           f"{per} per rank, no collective in the timed path"}


def local_shape(name, world, rank):
    """This rank's share of a workload: the (batch x kv-head) units are cut into contiguous slices
    (starflashattention_b200.multi_gpu.partition); a slice is either whole batch entries or a
    range of the kv-heads of one entry.  Returns the local problem shape and its share of the work."""
    from starflashattention_b200.multi_gpu import partition
    kind, dtype, B, Hq, Hkv, S, D, causal = WORKLOADS[name]
    mine = partition(B * Hkv, world, rank)
    n = len(mine)
    g = Hq // Hkv
    if n % Hkv == 0 and mine.start % Hkv == 0:
        b_loc, hkv_loc = n // Hkv, Hkv
    elif n > 0 and mine.start // Hkv == (mine.stop - 1) // Hkv:
        b_loc, hkv_loc = 1, n
    else:
        raise SystemExit(f"{name}: {B * Hkv} units do not cut into whole entries or head ranges over {world} ranks")
    return (kind, dtype, b_loc, hkv_loc * g, hkv_loc, S, D, causal), n / (B * Hkv)


def step_fn(sfa, shape, d, out):
    kind, dtype, B, Hq, Hkv, S, D, causal = shape
    if kind == "fwd":
        return lambda: sfa.fwd(d["q"], d["k"], d["v"], causal=causal, out=out, return_lse=True)
    return lambda: sfa.decode(d["q"], d["k"], d["v"], d["seqlens"], out=out)


def timed(fn, steps, dev, world, dist, flush=None, spin=0):
    """K calls between two CUDA events on the current stream (after a barrier + synchronise), max over ranks.
    spin: device-side cycles to burn ahead of the first event, so that the host has every call enqueued before the
    GPU reaches the timed region (calls of a few tens of microseconds would otherwise be timed at the host's pace)."""
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if flush is None:
        if spin:
            torch.cuda._sleep(int(spin))
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
    else:           # L2 flushed before every call: one event pair per call
        ms = 0.0
        wbuf, rbuf = flush
        for _ in range(steps):
            # 256 MB written (evicts everything), then 256 MB read: the written lines would otherwise still be
            # dirty in L2 and their write-back would share HBM with the timed kernel's reads
            wbuf.zero_()
            rbuf.max()
            # ~0.3 ms of device-side spinning: the host has enqueued the timed call before the GPU gets to e0, so
            # the interval holds the kernels and not the host's launch path
            torch.cuda._sleep(600000)
            e0.record(); fn(); e1.record()
            torch.cuda.synchronize()
            ms += e0.elapsed_time(e1)
    if world > 1:
        dist.barrier()
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()), ms


def copy_floor(host, out_host, lse_host, dev, world, dist, reps=3):
    """The box's own ceiling for the e2e leg: plain cudaMemcpyAsync of the step's input bytes (pinned host -> device)
    on one stream and of its output bytes (device -> pinned host) on another, at the same time, on every rank at once;
    no kernel.  Best of `reps`, max over ranks."""
    s_in, s_out = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
    d_in = {k_: torch.empty(v_.shape, dtype=v_.dtype, device=dev) for k_, v_ in host.items()}
    d_out, d_lse = torch.empty(out_host.shape, dtype=out_host.dtype, device=dev), torch.empty(lse_host.shape, dtype=torch.float32, device=dev)
    best = float("inf")
    for _ in range(reps):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        cur = torch.cuda.current_stream()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(cur)
        s_in.wait_event(e0); s_out.wait_event(e0)
        with torch.cuda.stream(s_in):
            for k_ in host:
                d_in[k_].copy_(host[k_], non_blocking=True)
        with torch.cuda.stream(s_out):
            out_host.copy_(d_out, non_blocking=True)
            lse_host.copy_(d_lse, non_blocking=True)
        cur.wait_stream(s_in); cur.wait_stream(s_out)
        e1.record(cur)
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    t = torch.tensor([best], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    del d_in, d_out, d_lse
    return float(t.item())


def source_hash(*names):
    import hashlib
    h = hashlib.sha256()
    for n in names:
        with open(os.path.join(ROOT, "starflashattention_b200", "csrc", n), "rb") as f:
            h.update(f.read())
    return h.hexdigest()[:16]


def load_traffic(name):
    """DRAM bytes per launch of the dominant kernel from the committed `ncu --set full` capture
    (profiles/r2_traffic.json).  The entry names the kernel and carries a hash of the kernel's source file; it is
    printed only while that hash still matches the source this run was built from."""
    try:
        e = json.load(open(os.path.join(ROOT, "profiles", "r2_traffic.json")))[name]
        if e.get("source_hash") != source_hash(*e["sources"]):
            return None
        return {"bytes_per_launch": float(e["bytes_per_launch"]), "kernel": e["kernel"], "capture": e.get("capture")}
    except Exception:  # noqa: BLE001
        return None


def run_ours(args):
    import starflashattention_b200 as sfa
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the library has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    peaks = load_peaks()

    name = args.workload
    kind, dtype, B, Hq, Hkv, S, D, causal = WORKLOADS[name]
    units = workload_units(name)
    # ONE problem per step, its (batch x kv-head) units cut over the ranks with no collective in the timed path
    # (strong scaling): C3 at N=8 is 32 units = one batch entry per GPU, C5 four heads per GPU, C4 one sequence per GPU.
    shape, share = local_shape(name, world, rank)
    _, _, Bl, Hql, Hkvl, _, _, _ = shape
    host, d = make_inputs(shape, dev, seed=100 + rank)
    out = torch.empty((Bl, S, Hql, D) if kind == "fwd" else (Bl, Hql, D), dtype=dtype, device=dev)
    step = step_fn(sfa, shape, d, out)

    # ---- device-resident timing
    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = sfa.launch_count()
    ms_max, ms_own = timed(step, args.steps, dev, world, dist)
    launches = sfa.launch_count() - launches0
    clocks = sampler.stop() if rank == 0 else None
    if kind == "fwd":
        wd = sfa.fwd_watchdog()
        if wd:
            raise SystemExit(f"prefill pipeline watchdog fired (code {wd}); the timing is void")
    ms_per_step = ms_max / args.steps
    value = units["amount"] / (ms_per_step / 1e3)          # the whole problem over the slowest rank's time

    # ---- end to end through the public host-buffer API (sfa_fwd_host / sfa_decode_host in the C ABI): pinned host
    # inputs -> H2D -> kernel -> D2H of the result, every step; the library cuts the call into pieces so that upload,
    # kernel and download overlap on three streams.
    out_host = torch.empty(out.shape, dtype=dtype, pin_memory=True)
    lse_host = torch.empty((Bl, Hql, S) if kind == "fwd" else (Bl, Hql), dtype=torch.float32, pin_memory=True)
    seq_host = host.get("seqlens")

    def e2e_step():
        if kind == "fwd":
            sfa.fwd_host(host["q"], host["k"], host["v"], causal=causal, out=out_host, lse=lse_host, device=dev)
        else:
            sfa.decode_host(host["q"], host["k"], host["v"], seq_host, out=out_host, lse=lse_host, device=dev)
    for _ in range(2):
        e2e_step()
    launches_e2e0 = sfa.launch_count()
    e2e_max, _ = timed(e2e_step, args.steps, dev, world, dist)
    launches_e2e = (sfa.launch_count() - launches_e2e0) // args.steps
    e2e_ms = e2e_max / args.steps
    h2d = sum(v.numel() * v.element_size() for v in host.values())
    d2h = out_host.numel() * out_host.element_size() + lse_host.numel() * lse_host.element_size()
    floor_ms = copy_floor(host, out_host, lse_host, dev, world, dist)

    # ---- a seconds-long back-to-back loop of the same step (what the clock settles at under the power cap)
    sustained = None
    if kind == "fwd" and world == 1 and not args.no_sustained:
        n_sus = max(args.steps, int(2500.0 / max(ms_per_step, 1e-3)))
        s2 = ClockSampler(local_rank)
        s2.start()
        sus_max, _ = timed(step, n_sus, dev, world, dist)
        c2 = s2.stop()
        rate = units["amount"] / (sus_max / n_sus / 1e3)
        sustained = {"value": rate, "unit": units["unit"], "steps": n_sus, "seconds": sus_max / 1e3, "ms_per_step": sus_max / n_sus,
                     "peak": peaks["tensor_sustained"], "frac": rate / peaks["tensor_sustained"], "clocks": c2,
                     "peak_source": peaks["source"] + " sustained (bf16_tflops_sustained)"}

    if sustained:
        time.sleep(2.0)      # let the power-capped clock recover before the secondary lines
    sec = None
    if not args.no_secondary:
        sec = secondary(sfa, name, dev, peaks, world, rank, dist)
    if rank == 0:
        peak = peaks[units["bound"]]
        per_gpu = units["amount"] * share / (ms_own / args.steps / 1e3)      # rank 0's own kernel rate on its share
        line = {
            "metric": units["metric"], "value": value, "unit": units["unit"], "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong",
            "vs_baseline": None, "dtype": "bf16" if dtype == torch.bfloat16 else "fp16", "data": "synthetic",
            "config": workload_config(name, world),
            "e2e": {"value": units["amount"] / (e2e_ms / 1e3), "unit": units["unit"],
                    "h2d_bytes_per_step": h2d * world, "d2h_bytes_per_step": d2h * world, "ms_per_step": e2e_ms,
                    "api": "sfa_fwd_host (C ABI) via fwd_host" if kind == "fwd" else "sfa_decode_host (C ABI) via decode_host",
                    "launches_per_step": int(launches_e2e), "steps": args.steps,
                    "copy_floor_ms": floor_ms, "over_copy_floor": e2e_ms / floor_ms if floor_ms > 0 else None,
                    "copy_floor_note": "plain pinned cudaMemcpyAsync of the same bytes, H2D and D2H streams at once, all ranks "
                                       "at once, no kernel; per-rank bytes = the totals / n_gpus"},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": units["bound"], "achieved": per_gpu, "peak": peak,
                         "unit": "TFLOP/s" if units["bound"] == "tensor" else "GB/s", "frac": per_gpu / peak,
                         "traffic": None, "algorithmic_bytes": algorithmic_bytes(name) * share,
                         "peak_source": peaks["source"] + (" burst (bf16_tflops)" if units["bound"] == "tensor" else " (hbm_gbs)"),
                         "per": "rank 0's share of the problem over rank 0's own event time"},
        }
        tr = load_traffic(name) if world == 1 else None
        if tr:
            line["roofline"]["traffic"] = tr["bytes_per_launch"]
            line["roofline"]["traffic_source"] = {"kernel": tr["kernel"], "capture": tr["capture"],
                                                  "note": "ncu --set full dram__bytes_read.sum + dram__bytes_write.sum per launch; "
                                                          "printed only while the kernel source hash matches the capture's"}
        if units["bound"] == "hbm":
            # MEASURED_PEAKS.json's hbm_gbs is a read + write copy; a read-only stream has no bus turnarounds and can
            # sit a few percent above it, so frac may exceed 1
            line["roofline"]["peak_note"] = "peak = measured read+write copy bandwidth; the decode stream is read-only"
        if sustained:
            line["sustained"] = sustained
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(name, budget_s=args.cpu_budget)
        if sec:
            line["secondary"] = sec
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def secondary(sfa, primary, dev, peaks, world, rank, dist):
    """Device-resident timings of the other BASELINE configs, sharded like the primary (explanatory, not the
    headline): each with its own roofline object and clock samples.  At N=1 also the section-8(d) decode variants."""
    res = {}
    # the power-hungry C5 goes last: the cases after it would run at the clocks the 1 kW cap leaves behind
    cases = [("c4", None)] + ([("c2", None), ("c4_b1_flushed", ("decode", torch.bfloat16, 1, 32, 8, 32768, 128, False)),
                               ("c4_b1_rotating", ("decode", torch.bfloat16, 1, 32, 8, 32768, 128, False)),
                               ("c4_b32", ("decode", torch.bfloat16, 32, 32, 8, 32768, 128, False))] if world == 1 else []) + [("c5", None)]
    for name, custom in cases:
        if name == primary:
            continue
        time.sleep(1.0)   # let the clocks recover from the case before (power cap): each line is measured from idle
        try:
            if custom is None:
                shape, share = local_shape(name, world, rank)
                full = WORKLOADS[name]
            else:
                shape, share, full = custom, 1.0, custom
            kind, dtype, B, Hq, Hkv, S, D, causal = full
            amount = (4.0 * B * Hq * S * S * D * (0.5 if causal else 1.0) / 1e12) if kind == "fwd" else (2.0 * B * S * Hkv * D * 2 / 1e9)
            bound = "tensor" if kind == "fwd" else "hbm"
            _, _, Bl, Hql, Hkvl, _, _, _ = shape
            g = torch.Generator(device=dev).manual_seed(7 + rank)
            mk = lambda shp: torch.randn(shp, generator=g, device=dev, dtype=torch.float32).to(dtype)
            if kind == "fwd":
                q, k, v = mk((Bl, S, Hql, D)), mk((Bl, S, Hkvl, D)), mk((Bl, S, Hkvl, D))
                fn = lambda: sfa.fwd(q, k, v, causal=causal)
            else:
                q, k, v = mk((Bl, Hql, D)), mk((Bl, S, Hkvl, D)), mk((Bl, S, Hkvl, D))
                sl = torch.full((Bl,), S, dtype=torch.int32, device=dev)
                fn = lambda: sfa.decode(q, k, v, sl)
            # inputs that fit the 126 MB L2 are flushed out before every call
            in_bytes = 2.0 * (q.numel() + k.numel() + v.numel())
            flush = ((torch.empty(256 << 20, dtype=torch.uint8, device=dev), torch.zeros(256 << 20, dtype=torch.uint8, device=dev))
                     if in_bytes < 400e6 else None)
            l2_note = None
            if name.endswith("_rotating"):
                # back-to-back calls over several caches that together exceed L2: every call streams cold data, and
                # the ~2 us granularity of a single event pair (the flushed line above) averages out over the calls
                flush = None
                n_sets = int(400e6 // in_bytes) + 2
                ks, vs = [k] + [torch.randn_like(k) for _ in range(n_sets - 1)], [v] + [torch.randn_like(v) for _ in range(n_sets - 1)]
                turn = [0]

                def fn():
                    i = turn[0] % n_sets
                    turn[0] += 1
                    return sfa.decode(q, ks[i], vs[i], sl, out=o_buf)
                o_buf = torch.empty_like(q)
                l2_note = f"{n_sets} caches of {in_bytes / 1e6:.0f} MB visited in turn by back-to-back calls (together larger than L2); all calls enqueued behind a device-side spin before the first event"
            for _ in range(3):
                fn()
            n = 20 if flush is not None else (5 if name == "c5" and world == 1 else (96 if l2_note else 20))
            smp = ClockSampler(dev.index)
            if rank == 0:
                smp.start()
            ms_max, ms_own = timed(fn, n, dev, world, dist, flush=flush, spin=8_000_000 if l2_note else 0)
            clk = smp.stop() if rank == 0 else None
            ms = ms_max / n
            rate = amount / (ms / 1e3)
            own = amount * share / (ms_own / n / 1e3)
            res[name] = {"metric": "attn_fwd_tflops" if kind == "fwd" else "decode_kv_gbs", "value": rate,
                         "unit": "TFLOP/s" if kind == "fwd" else "GB/s", "ms": ms, "steps": n,
                         "config": f"{kind} B={B} Hq={Hq} Hkv={Hkv} N={S} d={D} causal={causal}; {Bl} x {Hkvl} (batch x kv-head) per rank",
                         "l2": l2_note or ("flushed before every call (256 MB written, then 256 MB read so that no dirty line is left; one "
                                           "event pair per call, ~2 us granularity)" if flush is not None else "inputs larger than L2"),
                         "roofline": {"bound": bound, "achieved": own, "peak": peaks[bound], "frac": own / peaks[bound],
                                      "unit": "TFLOP/s" if kind == "fwd" else "GB/s"},
                         "clocks": clk}
            del q, k, v
            ks = vs = None
            torch.cuda.empty_cache()
        except Exception as e:  # noqa: BLE001 - a secondary line must not void the headline
            res[name] = {"error": repr(e)}
    return res


def cpu_baseline(name, budget_s=15.0):
    """The reference's CPU path (naive attention, oracle/naive.py) on a bounded sample of the workload."""
    from oracle import naive
    kind, dtype, B, Hq, Hkv, S, D, causal = WORKLOADS[name]
    units = workload_units(name)
    # all the host threads this process may use: torchrun exports OMP_NUM_THREADS=1 to its ranks, which would time the
    # reference's CPU path on one core at N > 1 (only rank 0 runs it, the other ranks do no CPU work)
    try:
        avail = len(os.sched_getaffinity(0))
    except AttributeError:
        avail = os.cpu_count() or 1
    if torch.get_num_threads() != avail:
        torch.set_num_threads(avail)
    cores = torch.get_num_threads()
    g = torch.Generator().manual_seed(0)
    t_used, n = 0.0, 0
    if kind == "fwd":
        per_slice = units["amount"] / (B * Hq)
        while t_used < budget_s and n < B * Hq:
            q, k, v = (torch.randn((1, S, 1, D), generator=g).to(dtype) for _ in range(3))
            t0 = time.perf_counter()
            naive.attention_fwd(q, k, v, causal=causal, chunk=2048)
            t_used += time.perf_counter() - t0
            n += 1
        sample = f"{n} of {B * Hq} (batch, head) slices of {name}, fp32 math on {dtype}-rounded inputs"
    else:
        per_slice = units["amount"] / B
        while t_used < budget_s and n < B:
            q = torch.randn((1, Hq, D), generator=g).to(dtype)
            k, v = (torch.randn((1, S, Hkv, D), generator=g).to(dtype) for _ in range(2))
            t0 = time.perf_counter()
            naive.decode_attention(q, k, v, torch.tensor([S]))
            t_used += time.perf_counter() - t0
            n += 1
        sample = f"{n} of {B} sequences of {name}, fp32 math on {dtype}-rounded inputs"
    return {"value": n * per_slice / t_used, "unit": units["unit"], "cores": cores, "kind": "port", "sample": sample,
            "host_cpus": os.cpu_count(), "seconds": t_used}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    name = args.workload
    units = workload_units(name)
    kind, dtype, B, Hq, Hkv, S, D, causal = WORKLOADS[name]
    per_step_budget = max(2.0, min(20.0, 120.0 / max(1, args.steps + args.warmup)))
    for _ in range(args.warmup):
        cpu_baseline(name, budget_s=per_step_budget)
    vals, secs = [], 0.0
    for _ in range(args.steps):
        r = cpu_baseline(name, budget_s=per_step_budget)
        vals.append(r["value"]); secs += r["seconds"]
    value = statistics.mean(vals)
    line = {
        "impl": "reference", "metric": units["metric"], "value": value, "unit": units["unit"],
        "n_gpus": int(os.environ.get("WORLD_SIZE", "1")), "steps": args.steps, "warmup": args.warmup,
        # time of ONE FULL step of the workload at the measured rate (each timed step is a bounded sample of it)
        "ms_per_step": 1e3 * units["amount"] / value, "sample_ms_per_step": 1e3 * secs / max(1, args.steps),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "fp32", "data": "synthetic",
        "config": workload_config(name, int(os.environ.get("WORLD_SIZE", "1"))),
        "note": "the reference ships no CPU kernel; its CPU path is the naive attention of "
                "examples/python/testFlashDecoder.py:84-92, restated in oracle/naive.py; rank 0 runs it on all host cores",
        "cpu_baseline": {"value": value, "unit": units["unit"], "cores": r["cores"], "kind": "port", "sample": r["sample"] + " per step"},
        "e2e": {"value": value, "unit": units["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--cpu-budget", type=float, default=15.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true")
    ap.add_argument("--no-sustained", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
