"""Decode with few (batch, kv-head) units and a long cache: time against the split count around one wave of CTAs
(148 SMs x 4 = 592 slots), behind a clean L2 flush.  The per-CTA time goes with the steps of 16 keys its busiest warp
makes: chunk = ceil(len / splits) rounded up to 64 keys (4 warps x 16), so 72 and 64 splits of 32768 keys both give
512-key chunks (8 steps per warp) while 74 give 448 (7 steps).

    python tools/decode_few_units.py
"""
import os
import statistics
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import starflashattention_b200 as sfa  # noqa: E402


def timed(fn, wbuf, rbuf, reps=25):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(reps):
        wbuf.zero_(); rbuf.sum(); torch.cuda._sleep(600000)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return statistics.median(ts), min(ts)


def rotating(fn_of_set, n_sets, calls=48, reps=7):
    """Average time of back-to-back calls that walk over `n_sets` different caches (together larger than L2): every
    call reads cold data, and the 2 us granularity of a single event pair (visible in the numbers above) averages out."""
    for i in range(n_sets):
        fn_of_set(i)
    ts = []
    for _ in range(reps):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(calls):
            fn_of_set(i % n_sets)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3 / calls)
    return statistics.median(ts), min(ts)


def main():
    dt, D = torch.bfloat16, 128
    g = torch.Generator(device="cuda").manual_seed(0)
    wbuf = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    rbuf = torch.ones(256 << 20, dtype=torch.uint8, device="cuda")
    for (B, Hq, Hkv, M, counts) in [(1, 32, 8, 32768, (64, 72, 73, 74, 80, 96, 128, 147, 148)),
                                    (1, 32, 8, 16384, (32, 37, 64, 72, 74)), (2, 32, 8, 32768, (32, 36, 37, 64, 74)),
                                    (1, 32, 8, 65536, (64, 72, 74, 128, 147)), (1, 8, 8, 32768, (64, 72, 74))]:
        q = torch.randn((B, Hq, D), generator=g, device="cuda", dtype=torch.float32).to(dt)
        kc = torch.randn((B, M, Hkv, D), generator=g, device="cuda", dtype=torch.float32).to(dt)
        vc = torch.randn((B, M, Hkv, D), generator=g, device="cuda", dtype=torch.float32).to(dt)
        out = torch.empty_like(q)
        floor = 2.0 * B * M * Hkv * D * 2 / 6543.4e3
        auto = sfa.decode_num_splits(B, M, Hkv)
        med, mn = timed(lambda: sfa.decode(q, kc, vc, None, out=out), wbuf, rbuf)
        print(f"[few_units] B={B} Hq={Hq} Hkv={Hkv} M={M} floor {floor:.1f} us  auto ({auto} splits) median {med:.1f} min {mn:.1f}", flush=True)
        for s in counts:
            med, mn = timed(lambda: sfa.decode(q, kc, vc, None, out=out, num_splits=s), wbuf, rbuf)
            chunk = ((M + s - 1) // s + 63) // 64 * 64
            print(f"[few_units]    splits {s:3d} (chunk {chunk}, {(M + chunk - 1) // chunk} busy): median {med:.1f} min {mn:.1f}", flush=True)
        # the same, as the average of 48 back-to-back calls over 6 different caches (6 x >= 67 MB > L2)
        n_sets = 6
        kcs = [kc] + [torch.randn_like(kc) for _ in range(n_sets - 1)]
        vcs = [vc] + [torch.randn_like(vc) for _ in range(n_sets - 1)]
        for s in (0,) + tuple(counts):
            med, mn = rotating(lambda i: sfa.decode(q, kcs[i], vcs[i], None, out=out, num_splits=s), n_sets)
            print(f"[few_units]    rotating caches, splits {s if s else 'auto':>4}: median {med:.2f} us per call  min {mn:.2f}", flush=True)
        del kcs, vcs


if __name__ == "__main__":
    main()
