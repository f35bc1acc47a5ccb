"""End-to-end prefill from pinned host buffers: the library's default piece plan against explicit uniform piece counts
(`sfa_fwd_host_args.pieces`), beside the plain-copy floor of the same bytes, the kernels of the plan alone, and the
timeline of the default plan (sfa_tuning_set("host_trace", 1) + sfa_fwd_host_trace).

    python tools/e2e_pieces.py c3|c5|c5r2 [steps]
"""
import ctypes as C
import os
import statistics
import sys
import time

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import starflashattention_b200 as sfa  # noqa: E402
from starflashattention_b200 import _capi  # noqa: E402

CFG = {"c3": (torch.bfloat16, 8, 32, 32, 8192, 128, True), "c5": (torch.bfloat16, 1, 32, 32, 65536, 128, True),
       "c5r2": (torch.bfloat16, 1, 16, 16, 65536, 128, True)}


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "c3"
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 8
    dt, B, Hq, Hkv, S, D, causal = CFG[name]
    g = torch.Generator().manual_seed(0)
    q = torch.randn((B, S, Hq, D), generator=g).to(dt).pin_memory()
    k = torch.randn((B, S, Hkv, D), generator=g).to(dt).pin_memory()
    v = torch.randn((B, S, Hkv, D), generator=g).to(dt).pin_memory()
    out = torch.empty_like(q).pin_memory()
    lse = torch.empty((B, Hq, S), dtype=torch.float32).pin_memory()
    dq, dk, dv, do = (torch.empty(t.shape, dtype=t.dtype, device="cuda") for t in (q, k, v, out))
    s2 = torch.cuda.Stream()

    def wall(fn):
        ts = []
        for i in range(steps + 2):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            fn()
            torch.cuda.synchronize()
            if i >= 2:
                ts.append((time.perf_counter() - t0) * 1e3)
        return statistics.median(ts), min(ts)

    def floor():
        with torch.cuda.stream(s2):
            out.copy_(do, non_blocking=True)
        dq.copy_(q, non_blocking=True); dk.copy_(k, non_blocking=True); dv.copy_(v, non_blocking=True)

    def up_only():
        dq.copy_(q, non_blocking=True); dk.copy_(k, non_blocking=True); dv.copy_(v, non_blocking=True)

    print(f"[e2e_pieces] {name} plain-copy floor (up || down) median {wall(floor)[0]:.2f} ms, upload alone {wall(up_only)[0]:.2f} ms", flush=True)
    anatomy(name, B, Hq, Hkv, S, D, causal, q, k, v, out, lse, dq, dk, dv, do, wall)
    for ks in (1, 2, 1, 2):
        sfa.tuning_set("host_kernel_streams", ks)
        med, mn = wall(lambda: sfa.fwd_host(q, k, v, causal=causal, out=out, lse=lse))
        print(f"[e2e_pieces] {name} default plan, kernels on {ks} stream(s): median {med:.2f} ms  min {mn:.2f} ms", flush=True)
    for pieces in (0, 8, 16, 32, 64):
        med, mn = wall(lambda: sfa.fwd_host(q, k, v, causal=causal, out=out, lse=lse, chunks=pieces or None))
        print(f"[e2e_pieces] {name} pieces={pieces or 'default'} median {med:.2f} ms  min {mn:.2f} ms", flush=True)


def anatomy(name, B, Hq, Hkv, S, D, causal, q, k, v, out, lse, dq, dk, dv, do, wall):
    """Where the time of the default plan goes: its kernels alone on device-resident slices (no copies), one full
    kernel, and the per-piece timeline of one call."""
    dq.copy_(q); dk.copy_(k); dv.copy_(v)
    plan = sfa.fwd_host_plan(B, S, S, Hq, Hkv, D, causal=causal)
    grp = Hq // Hkv

    def plan_kernels():
        for b, hk0, hkn, r0, r1, _ in plan:
            keys = r1 if causal else S
            sfa.fwd(dq[b:b + 1, r0:r1, hk0 * grp:(hk0 + hkn) * grp], dk[b:b + 1, :keys, hk0:hk0 + hkn],
                    dv[b:b + 1, :keys, hk0:hk0 + hkn], causal=causal, out=do[b:b + 1, r0:r1, hk0 * grp:(hk0 + hkn) * grp])

    print(f"[e2e_pieces] {name} kernels of the default plan ({len(plan)} pieces), device-resident: median {wall(plan_kernels)[0]:.2f} ms; "
          f"one full kernel {wall(lambda: sfa.fwd(dq, dk, dv, causal=causal, out=do))[0]:.2f} ms", flush=True)
    sfa.tuning_set("host_trace", 1)
    sfa.fwd_host(q, k, v, causal=causal, out=out, lse=lse)
    buf = (C.c_float * (5 * 256))()
    n = _capi.lib.sfa_fwd_host_trace(buf, 256)
    sfa.tuning_set("host_trace", 0)
    for i in range(n):
        b, hk0, hkn, r0, r1, _ = plan[i]
        t = [buf[5 * i + j] for j in range(5)]
        print(f"[e2e_pieces] trace piece {i:2d} entry {b} heads {hk0:2d}+{hkn:<2d} rows {r0:6d}-{r1:6d}: up {t[0]:6.2f}-{t[1]:6.2f}  "
              f"kernel {t[2]:6.2f}-{t[3]:6.2f}  down -{t[4]:6.2f}", flush=True)


if __name__ == "__main__":
    main()
