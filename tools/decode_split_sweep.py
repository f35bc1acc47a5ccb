"""Split count and merge form of the split-KV decode: the two-launch form (split kernel + combine kernel) against the
in-cluster merge (splits <= 8 as one thread-block cluster, merged through distributed shared memory), over the
shapes between "launch bound" and "HBM bound".  One call per CUDA-event pair, 256 MB L2 flush before each, median
of 15.

    python tools/decode_split_sweep.py [quick]
"""
import json
import os
import statistics
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import starflashattention_b200 as sfa  # noqa: E402

SHAPES = [  # B, Hq, Hkv, M
    (1, 32, 32, 512), (2, 32, 32, 4096), (2, 32, 32, 8192), (1, 32, 8, 4096), (8, 32, 8, 4096), (4, 32, 8, 4096), (4, 32, 8, 16384),
    (32, 32, 8, 2048), (1, 32, 8, 32768), (4, 32, 8, 32768), (8, 32, 8, 32768),
]


def timed(fn, flush, reps=15):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return statistics.median(ts)


def main():
    dt, D = torch.bfloat16, 128
    g = torch.Generator(device="cuda").manual_seed(0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    out = {}
    for (B, Hq, Hkv, M) in SHAPES:
        q = torch.randn((B, Hq, D), generator=g, device="cuda", dtype=torch.float32).to(dt)
        kc = torch.randn((B, M, Hkv, D), generator=g, device="cuda", dtype=torch.float32).to(dt)
        vc = torch.randn((B, M, Hkv, D), generator=g, device="cuda", dtype=torch.float32).to(dt)
        sl = torch.full((B,), M, dtype=torch.int32, device="cuda")
        nbytes = 2.0 * B * M * Hkv * D * 2
        auto = sfa.decode_num_splits(B, M, Hkv)
        row = {"auto_splits": auto, "floor_us": nbytes / 6556.8e3}
        sfa.tuning_set("decode_cluster", 1)
        row["auto"] = timed(lambda: sfa.decode(q, kc, vc, sl), flush)
        for s in (1, 2, 4, 6, 8, 9, 10, 12, 14, 16, 18, 24, 32, 48):
            if s * 64 > M and s > 1:
                continue
            for cl in ("0", "1"):
                if cl == "1" and not (2 <= s <= 8):
                    continue
                sfa.tuning_set("decode_cluster", int(cl))
                row[f"s{s}" + ("c" if cl == "1" else "")] = timed(lambda: sfa.decode(q, kc, vc, sl, num_splits=s), flush)
        sfa.tuning_set("decode_cluster", 1)
        out[f"B{B}_Hq{Hq}_Hkv{Hkv}_M{M}"] = row
        best = min((v, k) for k, v in row.items() if k.startswith("s"))
        print(f"[split_sweep] B={B} Hq={Hq} Hkv={Hkv} M={M}: floor {row['floor_us']:.1f} us, auto (S={auto}) {row['auto']:.1f} us, "
              f"best {best[1]} {best[0]:.1f} us | " +
              " ".join(f"{k}={v:.1f}" for k, v in row.items() if k.startswith("s")), flush=True)
        del q, kc, vc
    print(json.dumps(out))


if __name__ == "__main__":
    main()
