"""Does the split-KV decode kernel run faster when every head's rows are contiguous?  Same kernel, same bytes, two
cache layouts through strides: (B, M, Hkv, D) - the documented one, 256-byte pieces per head - and head-major
(B, Hkv, M, D)."""
import os, statistics, sys
import torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import starflashattention_b200 as sfa

B, Hq, Hkv, M, D = 8, 32, 8, 32768, 128
dt = torch.bfloat16
g = torch.Generator(device="cuda").manual_seed(0)
q = torch.randn((B, Hq, D), generator=g, device="cuda").to(dt)
sl = torch.full((B,), M, dtype=torch.int32, device="cuda")
for name in ("row-major (B,M,Hkv,D)", "head-major (B,Hkv,M,D)"):
    if name.startswith("row"):
        k = torch.randn((B, M, Hkv, D), generator=g, device="cuda").to(dt); v = torch.randn((B, M, Hkv, D), generator=g, device="cuda").to(dt)
    else:
        k = torch.randn((B, Hkv, M, D), generator=g, device="cuda").to(dt).transpose(1, 2)
        v = torch.randn((B, Hkv, M, D), generator=g, device="cuda").to(dt).transpose(1, 2)
    for _ in range(3):
        sfa.decode(q, k, v, sl)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(30):
        sfa.decode(q, k, v, sl)
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 30 * 1e3
    print(f"[layout] {name}: {us:.1f} us  {2.0 * B * M * Hkv * D * 2 / us / 1e3:.0f} GB/s", flush=True)
    del k, v
