"""Device-resident timing of the prefill kernel for one configuration (env overrides SFA_FWD_POLY / SFA_FWD_DBG
are read by the library at first launch, so every variant runs in its own process).

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
    if os.environ.get("SFA_FWD_DBG") in ("4", "5"):
        import ctypes
        from starflashattention_b200 import _capi
        buf = (ctypes.c_uint64 * 32)()
        _capi.lib.sfa_fwd_probe_read(buf, 32)
        v_ = list(buf)
        for qq in (0, 1):
            n = max(1, v_[8 * qq + 3])
            print(f"[probe] softmax q{qq}: tiles {n}  per tile: wait S {v_[8*qq]/n:.0f}  ld+max {v_[8*qq+1]/n:.0f}  "
                  f"exp+store {v_[8*qq+2]/n:.0f} (of which wait PV {v_[8*qq+4]/n:.0f}) cycles")
        n = max(1, v_[20])
        print(f"[probe] mma warp: iters {n}  per iter: wait P {v_[16]/n:.0f}  wait S-free {v_[17]/n:.0f}  wait KV {v_[18]/n:.0f}  total {v_[19]/n:.0f} cycles")
    print(f"[fwd_time] {name} POLY={os.environ.get('SFA_FWD_POLY','-')} DBG={os.environ.get('SFA_FWD_DBG','-')} STAGES={os.environ.get('SFA_FWD_STAGES','-')} "
          f"median {ms:.3f} ms  min {min(ts):.3f} ms  {flops / ms / 1e9:.1f} TFLOP/s (median)  watchdog={wd}{clk}", flush=True)


if __name__ == "__main__":
    main()
