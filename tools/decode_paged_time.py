"""Timing of the paged split-KV decode against the dense kernel at the C4 shape (B=8, 32/8 heads, KV 32768, bf16):
the same bytes through a shuffled block table, page sizes 16 .. 1024.  10 back-to-back steps per CUDA-event pair
(1 GiB of K/V per step, far beyond L2), median of 7.

    python tools/decode_paged_time.py [batch] [kv_len]
"""
import json
import os
import statistics
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import starflashattention_b200 as sfa  # noqa: E402


def timed(fn, reps=7, inner=10):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(inner):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3 / inner)
    return statistics.median(ts)


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    M = int(sys.argv[2]) if len(sys.argv) > 2 else 32768
    Hq, Hkv, D, dt = 32, 8, 128, torch.bfloat16
    g = torch.Generator(device="cuda").manual_seed(0)
    q = torch.randn((B, Hq, D), generator=g, device="cuda", dtype=torch.float32).to(dt)
    kc = torch.randn((B, M, Hkv, D), generator=g, device="cuda", dtype=torch.float32).to(dt)
    vc = torch.randn((B, M, Hkv, D), generator=g, device="cuda", dtype=torch.float32).to(dt)
    sl = torch.full((B,), M, dtype=torch.int32, device="cuda")
    nbytes = 2.0 * B * M * Hkv * D * 2
    out = {"shape": f"B={B} Hq={Hq} Hkv={Hkv} M={M} d={D} bf16", "bytes_per_step": nbytes}
    us = timed(lambda: sfa.decode(q, kc, vc, sl))
    o_d, lse_d = sfa.decode(q, kc, vc, sl)
    out["dense"] = {"us": us, "gbs": nbytes / us / 1e3}
    print(f"[decode_paged] dense: {us:.1f} us  {nbytes / us / 1e3:.0f} GB/s", flush=True)
    for page in (16, 64, 256, 1024):
        blocks = M // page
        # the pools ARE the dense caches cut into pages; the table scatters every sequence over the whole pool
        perm = torch.randperm(B * blocks, generator=g, device="cuda")
        k_pool = torch.empty((B * blocks, page, Hkv, D), dtype=dt, device="cuda")
        v_pool = torch.empty_like(k_pool)
        k_pool[perm] = kc.reshape(B * blocks, page, Hkv, D)
        v_pool[perm] = vc.reshape(B * blocks, page, Hkv, D)
        table = perm.reshape(B, blocks).to(torch.int32)
        us = timed(lambda: sfa.decode_paged(q, k_pool, v_pool, table, sl))
        o_p, lse_p = sfa.decode_paged(q, k_pool, v_pool, table, sl)
        same = bool(torch.equal(o_p, o_d) and torch.equal(lse_p, lse_d))
        out[f"page{page}"] = {"us": us, "gbs": nbytes / us / 1e3, "bit_identical_to_dense": same}
        print(f"[decode_paged] page {page}: {us:.1f} us  {nbytes / us / 1e3:.0f} GB/s  bit-identical to dense: {same}", flush=True)
        del k_pool, v_pool
    print(json.dumps(out))


if __name__ == "__main__":
    main()
