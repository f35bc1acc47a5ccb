"""Device-resident timing of the prefill kernel for one configuration (the SFA_FWD_* overrides are read by the
library once per process, so every variant runs in its own process; most act in the probe build only:
SFA_PROBES_LIB=1 after `python starflashattention_b200/build.py --probes`).

    python tools/fwd_time.py c3|c2|c5 [iters]
"""
import os
import statistics
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import starflashattention_b200 as sfa  # noqa: E402

CFG = {"c3": (torch.bfloat16, 8, 32, 8192, 128, True), "c2": (torch.float16, 4, 16, 4096, 64, False),
       "c5": (torch.bfloat16, 1, 32, 65536, 128, True), "c3nc": (torch.bfloat16, 8, 32, 8192, 128, False),
       "c2b": (torch.float16, 16, 16, 4096, 64, False),
       # one rank's share of C3 / C5 on 8 GPUs
       "c3r8": (torch.bfloat16, 1, 32, 8192, 128, True), "c5r8": (torch.bfloat16, 1, 4, 65536, 128, True),
       # head_dim 64 shapes for the warpgroup-stagger sweep
       "d64c": (torch.bfloat16, 4, 16, 4096, 64, True), "d64s": (torch.float16, 16, 16, 1024, 64, False),
       "d64l": (torch.bfloat16, 1, 16, 16384, 64, True)}


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "c3"
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    dt, B, H, S, D, causal = CFG[name]
    g = torch.Generator(device="cuda").manual_seed(0)
    q, k, v = (torch.randn((B, S, H, D), generator=g, device="cuda", dtype=torch.float32).to(dt) for _ in range(3))
    out = torch.empty_like(q)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(3):
        sfa.fwd(q, k, v, causal=causal, out=out)
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); sfa.fwd(q, k, v, causal=causal, out=out); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = statistics.median(ts)
    flops = 4.0 * B * H * S * S * D * (0.5 if causal else 1.0)
    wd = sfa.fwd_watchdog()
    # sustained loop with NVML clock sampling (what the SM clock does under a seconds-long tensor load)
    clk = ""
    try:
        import threading, time, pynvml as nv
        nv.nvmlInit(); h = nv.nvmlDeviceGetHandleByIndex(0)
        stop, sm, pw = threading.Event(), [], []
        def loop():
            while not stop.is_set():
                sm.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)); pw.append(nv.nvmlDeviceGetPowerUsage(h) / 1e3)
                time.sleep(0.01)
        th = threading.Thread(target=loop); th.start()
        n_sus = max(20, int(1000 / ms))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n_sus):
            sfa.fwd(q, k, v, causal=causal, out=out)
        e1.record(); torch.cuda.synchronize()
        stop.set(); th.join()
        half = sm[len(sm) // 2:]
        clk = (f" | sustained x{n_sus}: {e0.elapsed_time(e1) / n_sus:.3f} ms/iter, SM clock median(2nd half) "
               f"{statistics.median(half):.0f} MHz min {min(sm)} max {max(sm)}, power max {max(pw):.0f} W")
    except Exception as e:  # noqa: BLE001
        clk = f" | clock sampling failed: {e!r}"
    tag = " ".join(f"{k[8:]}={v}" for k, v in sorted(os.environ.items()) if k.startswith("SFA_FWD_")) or "product"
    print(f"[fwd_time] {name} [{tag}{' probes-lib' if os.environ.get('SFA_PROBES_LIB') == '1' else ''}] "
          f"median {ms:.3f} ms  min {min(ts):.3f} ms  {flops / ms / 1e9:.1f} TFLOP/s (median)  watchdog={wd}{clk}", flush=True)


if __name__ == "__main__":
    main()
