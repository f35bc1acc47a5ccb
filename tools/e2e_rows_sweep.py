"""sfa_fwd_host on C5-like shapes (one long causal sequence, all heads / a rank's half / a rank's eighth) against the
explicit piece count: with 4-head groups, `pieces` sets the rows per piece (profiles/r2_ks_sweep.txt; the one-wave
rule of the heavy plan in csrc/host_pipeline.cu comes from this sweep).

    python tools/e2e_rows_sweep.py
"""
import os
import statistics
import sys
import time

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import starflashattention_b200 as sfa  # noqa: E402


def run(name, B, H, S, D=128):
    dt = torch.bfloat16
    g = torch.Generator().manual_seed(0)
    q, k, v = (torch.randn((B, S, H, D), generator=g).to(dt).pin_memory() for _ in range(3))
    out = torch.empty_like(q).pin_memory()
    lse = torch.empty((B, H, S), dtype=torch.float32).pin_memory()

    def wall(fn, steps=8):
        ts = []
        for i in range(steps + 2):
            torch.cuda.synchronize(); t0 = time.perf_counter(); fn(); torch.cuda.synchronize()
            if i >= 2:
                ts.append((time.perf_counter() - t0) * 1e3)
        return statistics.median(ts), min(ts)

    for _ in range(2):
        for pieces in (0, 16, 32, 48, 64, 96, 128):
            med, mn = wall(lambda: sfa.fwd_host(q, k, v, causal=True, out=out, lse=lse, chunks=pieces or None))
            print(f"[ks_sweep] {name} pieces={pieces or 'default'} ({sfa.fwd_host_last_pieces()}) median {med:.2f} min {mn:.2f}", flush=True)


if __name__ == "__main__":
    run("c5", 1, 32, 65536)
    run("c5r2", 1, 16, 65536)
    run("c5r8", 1, 4, 65536)
