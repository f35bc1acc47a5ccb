"""Event trace of one CTA of the prefill kernel (SFA_FWD_DBG=6): where the MMA warp and the softmax warpgroups
wait for each other.  Prints clock64() times relative to the first event of the window.

    SFA_FWD_DBG=6 python tools/fwd_trace.py [first_tile] [tiles]
"""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
os.environ.setdefault("SFA_FWD_DBG", "6")
import starflashattention_b200 as sfa  # noqa: E402
from starflashattention_b200 import _capi  # noqa: E402

NAMES = {7: "mma: QK step start", 0: "mma: K tile and S buffer acquired", 1: "mma: commits flushed, QK issue",
         13: "mma: QK 8 MMAs accepted", 14: "mma: V tile acquired", 2: "mma: P half 0 acquired",
         3: "mma: commits flushed, PVa issue", 4: "mma: PVa issued", 5: "mma: P half 1 acquired, PVb issue",
         15: "mma: PVb 4 MMAs accepted",
         8: "wg : S seen", 9: "wg : S-free arrive", 10: "wg : exp start", 11: "wg : P half 0 arrive",
         12: "wg : P half 1 arrive"}


def main():
    j0 = int(sys.argv[1]) if len(sys.argv) > 1 else 30
    nj = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    B, H, S, D = 2, 32, 8192, 128
    g = torch.Generator(device="cuda").manual_seed(0)
    q, k, v = (torch.randn((B, S, H, D), generator=g, device="cuda").to(torch.bfloat16) for _ in range(3))
    for _ in range(2):
        sfa.fwd(q, k, v, causal=True)
    torch.cuda.synchronize()
    n = 32 + 2048
    buf = (ctypes.c_uint64 * n)()
    _capi.lib.sfa_fwd_probe_read(buf, n)
    ev = []
    for j in range(j0, j0 + nj):
        for qq in (0, 1):
            for e, name in NAMES.items():
                t = buf[32 + (j * 2 + qq) * 16 + e]
                if t:
                    ev.append((t, f"q{qq} j{j} {name}"))
    ev.sort()
    t0 = ev[0][0]
    for t, name in ev:
        print(f"{t - t0:7d}  {name}")
    # prologue / epilogue of the CTA (the heaviest query block: 64 key tiles per query tile)
    t_in, t_setup, t_lastpv, t_stored, t_out = (buf[i] for i in (24, 25, 26, 27, 28))
    first_qk = buf[32 + 1]                          # q0 j0 "commits flushed, QK issue"
    first_s = buf[32 + 8]                           # q0 j0 "S seen"
    if t_in and t_out:
        print(f"CTA (0,0,0): {t_out - t_in} cycles in all; entry -> set-up done {t_setup - t_in}, -> first QK issue "
              f"{first_qk - t_in}, -> first S seen {first_s - t_in}; last P.V seen -> O stored {t_stored - t_lastpv}, "
              f"-> exit {t_out - t_lastpv}")
    # per-tile periods
    for qq in (0, 1):
        ts = [buf[32 + (j * 2 + qq) * 16 + 8] for j in range(8, 60)]
        d = [b - a for a, b in zip(ts, ts[1:])]
        print(f"q{qq}: mean cycles between consecutive 'S seen' (tiles 8..59): {sum(d) / len(d):.0f}")


if __name__ == "__main__":
    main()
