"""Event trace of CTA 0 of the persistent prefill kernel (probe build: python starflashattention_b200/build.py --probes):
where the MMA warps and the softmax warpgroups wait for each other, in steady state and across an item boundary.
Prints clock64() times relative to the first event of the window; tiles are numbered by the running count of key
tiles the CTA has processed for a query tile (CTA 0 starts with the heaviest item: 63 / 64 tiles at N = 8192).

    SFA_PROBES_LIB=1 python tools/fwd_trace.py [first_tile] [tiles]
"""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
os.environ["SFA_PROBES_LIB"] = "1"
import starflashattention_b200 as sfa  # noqa: E402
from starflashattention_b200 import _capi  # noqa: E402

NAMES = {0: "qk : K (and Q) tile acquired", 1: "qk : S buffer acquired, QK issue",
         2: "qk : QK issued + committed", 3: "pv : V tile, P half 0 acquired, PVa issue", 5: "pv : P half 1 acquired, PVb issue",
         6: "pv : PVb issued + committed",
         15: "wg : item start", 8: "wg : S seen", 9: "wg : S-free arrive", 10: "wg : exp start", 11: "wg : P half 0 arrive",
         12: "wg : P half 1 arrive", 13: "wg : epilogue of the item starts (last P.V seen)", 14: "wg : O stored (TMA store issued), LSE written"}


def main():
    j0 = int(sys.argv[1]) if len(sys.argv) > 1 else 30
    nj = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    B, H, S, D = 2, 32, 8192, 128
    g = torch.Generator(device="cuda").manual_seed(0)
    q, k, v = (torch.randn((B, S, H, D), generator=g, device="cuda").to(torch.bfloat16) for _ in range(3))
    for _ in range(2):
        sfa.fwd(q, k, v, causal=True)
    torch.cuda.synchronize()
    n = 32 + 4096
    buf = (ctypes.c_uint64 * n)()
    _capi.check(_capi.lib.sfa_fwd_probe_read(buf, n))

    def window(lo, hi, title):
        ev = []
        for j in range(lo, hi):
            for qq in (0, 1):
                for e, name in NAMES.items():
                    t = buf[32 + (j * 2 + qq) * 16 + e]
                    if t:
                        ev.append((t, f"q{qq} t{j} {name}"))
        ev.sort()
        if not ev:
            return
        print(f"---- {title}")
        t0 = ev[0][0]
        for t, name in ev:
            print(f"{t - t0:7d}  {name}")
    window(j0, j0 + nj, f"steady state, tiles {j0} .. {j0 + nj - 1}")
    window(61, 67, "item boundary of CTA 0 (first item: 63 / 64 tiles)")
    e = [buf[i] for i in range(11)]
    if e[10] and e[9]:
        print("last epilogue of CTA 0's warp 0 (cycles from its start): P.V seen +%d | half 0: O loaded +%d, staging free +%d, "
              "written + fenced + store issued +%d | half 1: start +%d, O loaded +%d, staging free +%d, store issued +%d | "
              "LSE written, done +%d" % (e[0] - e[9], e[1] - e[9], e[2] - e[9], e[3] - e[9], e[4] - e[9], e[5] - e[9], e[6] - e[9], e[7] - e[9], e[10] - e[9]))
    t_in, t_setup, t_out = (buf[i] for i in (24, 25, 28))
    if t_in and t_out:
        print(f"CTA 0: {t_out - t_in} cycles in all; entry -> set-up done {t_setup - t_in}")
    for qq in (0, 1):
        ts = [buf[32 + (j * 2 + qq) * 16 + 8] for j in range(8, 60)]
        d = [b - a for a, b in zip(ts, ts[1:])]
        print(f"q{qq}: mean cycles between consecutive 'S seen' (tiles 8..59): {sum(d) / len(d):.0f}")
        ts = [buf[32 + (j * 2 + qq) * 16 + 8] for j in range(70, 120)]
        d = [b - a for a, b in zip(ts, ts[1:]) if a and b]
        if d:
            print(f"q{qq}: mean cycles between consecutive 'S seen' (tiles 70..119, later items): {sum(d) / len(d):.0f}")


if __name__ == "__main__":
    main()
