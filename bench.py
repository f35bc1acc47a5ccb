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
independent, so every rank runs the full per-GPU workload with no collective in the timed path (weak scaling).
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


def load_traffic(name):
    """DRAM bytes per launch of the dominant kernel from the committed ncu capture (profiles/r1d_traffic.json)."""
    try:
        return float(json.load(open(os.path.join(ROOT, "profiles", "r1d_traffic.json")))[name]["bytes_per_launch"])
    except Exception:  # noqa: BLE001
        return None


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return {"hbm": float(p["hbm_gbs"]), "tensor": float(p["bf16_tflops"]), "source": "measured"}
    return {"hbm": 6650.0, "tensor": 1590.0, "source": "fallback"}


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


def make_inputs(name, device, seed):
    """Synthetic N(0,1) inputs: drawn on the device, mirrored into pinned host buffers for the e2e leg."""
    kind, dtype, B, Hq, Hkv, S, D, causal = WORKLOADS[name]
    g = torch.Generator(device=device).manual_seed(seed)

    def rnd(shape):
        t = torch.empty(shape, dtype=dtype, device=device)
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

    name = args.workload
    kind, dtype, B, Hq, Hkv, S, D, causal = WORKLOADS[name]
    units = workload_units(name)
    # c5 (BASELINE config 5) is ONE long-context problem whose heads are sharded over the ranks (strong scaling,
    # starflashattention_b200.multi_gpu.partition); every other workload runs in full on every rank (weak scaling).
    strong = name == "c5" and world > 1
    Hq_all, Hkv_all = Hq, Hkv
    if strong:
        my_heads = sfa.partition(Hkv, world, rank)
        frac = len(my_heads) / Hkv
        Hq, Hkv = Hq // Hkv * len(my_heads), len(my_heads)
        WORKLOADS[name] = (kind, dtype, B, Hq, Hkv, S, D, causal)
        units = dict(units, amount=units["amount"] * frac)
    host, d = make_inputs(name, dev, seed=100 + rank)
    if kind == "fwd":
        out = torch.empty((B, S, Hq, D), dtype=dtype, device=dev)

        def step():
            return sfa.fwd(d["q"], d["k"], d["v"], causal=causal, out=out)
    else:
        out = torch.empty((B, Hq, D), dtype=dtype, device=dev)

        def step():
            return sfa.decode(d["q"], d["k"], d["v"], d["seqlens"], out=out)
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing
    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = sfa.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    launches = sfa.launch_count() - launches0
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None
    if kind == "fwd":
        wd = sfa.fwd_watchdog()
        if wd:
            raise SystemExit(f"prefill pipeline watchdog fired (code {wd}); the timing is void")
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    ms_per_step = ms_max / args.steps
    total_amount = units["amount"]
    if world > 1:
        ta = torch.tensor([units["amount"]], dtype=torch.float64, device=dev)
        dist.all_reduce(ta, op=dist.ReduceOp.SUM)          # units all ranks processed (outside the timed region)
        total_amount = float(ta.item())
    value = total_amount / (ms_per_step / 1e3)

    # ---- end to end through the public host-buffer API (fwd_host / decode_host): pinned host inputs -> H2D ->
    # kernel -> D2H of the result, every step; the API cuts the independent (batch x head) units into chunks so
    # that upload, kernel and download overlap on three streams.
    out_host = torch.empty(out.shape, dtype=dtype, pin_memory=True)
    lse_shape = (B, Hq, S) if kind == "fwd" else (B, Hq)
    lse_host = torch.empty(lse_shape, dtype=torch.float32, pin_memory=True)
    seq_host = host.get("seqlens")
    e2e_steps = max(2, min(args.steps, 5))

    def e2e_step():
        if kind == "fwd":
            sfa.fwd_host(host["q"], host["k"], host["v"], causal=causal, out=out_host, lse=lse_host, device=dev)
        else:
            sfa.decode_host(host["q"], host["k"], host["v"], seq_host, out=out_host, lse=lse_host, device=dev)
    e2e_step()
    barrier()
    launches_e2e0 = sfa.launch_count()
    e0.record()
    for _ in range(e2e_steps):
        e2e_step()
    e1.record()
    barrier()
    launches_e2e = (sfa.launch_count() - launches_e2e0) // e2e_steps
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_ms = float(t.item()) / e2e_steps
    h2d = sum(v.numel() * v.element_size() for v in host.values())
    d2h = out_host.numel() * out_host.element_size() + lse_host.numel() * lse_host.element_size()

    if rank == 0:
        peaks = load_peaks()
        peak = peaks[units["bound"]]
        per_gpu = units["amount"] / (ms / args.steps / 1e3)      # rank 0's own kernel rate
        line = {
            "metric": units["metric"], "value": value, "unit": units["unit"], "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong" if strong else "weak",
            "vs_baseline": None, "dtype": "bf16" if dtype == torch.bfloat16 else "fp16", "data": "synthetic",
            "config": {"workload": f"{name}: {kind} B={B} Hq={Hq_all} Hkv={Hkv_all} N={S} d={D} causal={causal}",
                       "l2": "inputs larger than L2 (no flush)" if name != "c2" else "inputs 101 MB, L2-resident between steps",
                       "parallelism": (f"heads of one problem sharded over {world} rank(s), no collective" if strong else
                                       f"batch x heads units, {world} independent rank(s) each with the full per-GPU workload, no collective")},
            "e2e": {"value": total_amount / (e2e_ms / 1e3), "unit": units["unit"],
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms,
                    "api": "fwd_host" if kind == "fwd" else "decode_host", "launches_per_step": int(launches_e2e)},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": units["bound"], "achieved": per_gpu, "peak": peak,
                         "unit": "TFLOP/s" if units["bound"] == "tensor" else "GB/s", "frac": per_gpu / peak,
                         "traffic": load_traffic(name) if not strong else None, "traffic_unit": "bytes per launch (ncu)",
                         "algorithmic_bytes": algorithmic_bytes(name), "peak_source": peaks["source"] + (" burst" if units["bound"] == "tensor" else "")},
        }
        if units["bound"] == "hbm":
            # MEASURED_PEAKS.json's hbm_gbs is a read + write copy; a read-only stream has no bus turnarounds and can
            # sit a few percent above it, so frac may exceed 1
            line["roofline"]["peak_note"] = "peak = measured read+write copy bandwidth; the decode stream is read-only"
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(name, budget_s=args.cpu_budget)
        if world == 1 and not args.no_secondary:
            line["secondary"] = secondary(sfa, name, dev, peaks)
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def secondary(sfa, primary, dev, peaks):
    """Short device-resident timings of the other BASELINE configs (explanatory, not the headline)."""
    res = {}
    for name in ("c4", "c2"):
        if name == primary:
            continue
        try:
            kind, dtype, B, Hq, Hkv, S, D, causal = WORKLOADS[name]
            units = workload_units(name)
            g = torch.Generator(device=dev).manual_seed(7)
            mk = lambda shape: torch.randn(shape, generator=g, device=dev, dtype=torch.float32).to(dtype)
            if kind == "fwd":
                q, k, v = mk((B, S, Hq, D)), mk((B, S, Hkv, D)), mk((B, S, Hkv, D))
                fn = lambda: sfa.fwd(q, k, v, causal=causal)
            else:
                q, k, v = mk((B, Hq, D)), mk((B, S, Hkv, D)), mk((B, S, Hkv, D))
                sl = torch.full((B,), S, dtype=torch.int32, device=dev)
                fn = lambda: sfa.decode(q, k, v, sl)
            flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev) if name == "c2" else None
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            times = []
            # flushed workloads: one call per event pair behind the flush kernel (which also hides the host's
            # enqueue time); the others: 10 back-to-back calls per event pair, as the primary line is timed
            inner = 1 if flush is not None else 10
            for _ in range(10 if flush is not None else 5):
                if flush is not None:
                    flush.zero_()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(inner):
                    fn()
                e1.record()
                torch.cuda.synchronize()
                times.append(e0.elapsed_time(e1) / inner)
            ms = statistics.median(times)
            rate = units["amount"] / (ms / 1e3)
            res[name] = {"metric": units["metric"], "value": rate, "unit": units["unit"], "ms": ms,
                         "frac_of_measured_peak": rate / peaks[units["bound"]],
                         "l2": "flushed between iterations" if flush is not None else "inputs larger than L2"}
            del q, k, v
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
        "ms_per_step": 1e3 * secs / max(1, args.steps), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "fp32", "data": "synthetic",
        "config": {"workload": f"{name}: {kind} B={B} Hq={Hq} Hkv={Hkv} N={S} d={D} causal={causal}",
                   "note": "the reference ships no CPU kernel; its CPU path is the naive attention of "
                           "examples/python/testFlashDecoder.py:84-92, restated in oracle/naive.py"},
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
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
