"""Host time of one call of the public API (Python + C ABI + launch), measured on problems whose kernels are much
shorter than the call: wall clock over 2000 back-to-back calls, and the same with the GPU idle between calls."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import starflashattention_b200 as sfa  # noqa: E402

g = torch.Generator(device="cuda").manual_seed(0)
rnd = lambda shape, dt: torch.randn(shape, generator=g, device="cuda", dtype=torch.float32).to(dt)
q, k, v = (rnd((1, 256, 2, 128), torch.bfloat16) for _ in range(3))
out = torch.empty_like(q)
for name, fn in (("fwd(out=, lse)", lambda: sfa.fwd(q, k, v, causal=True, out=out)),
                 ("fwd(out=, no lse)", lambda: sfa.fwd(q, k, v, causal=True, out=out, return_lse=False))):
    for _ in range(50):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(2000):
        fn()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print(f"[host] {name}: {1e6 * (t1 - t0) / 2000:.1f} us per call to enqueue, {1e6 * (t2 - t0) / 2000:.1f} us per call until the GPU is done")
qd, kd, vd = rnd((4, 32, 128), torch.bfloat16), rnd((4, 1024, 8, 128), torch.bfloat16), rnd((4, 1024, 8, 128), torch.bfloat16)
sl = torch.full((4,), 1024, dtype=torch.int32, device="cuda")
od = torch.empty_like(qd)
for _ in range(50):
    sfa.decode(qd, kd, vd, sl, out=od)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(2000):
    sfa.decode(qd, kd, vd, sl, out=od)
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f"[host] decode(out=): {1e6 * (t1 - t0) / 2000:.1f} us per call to enqueue, {1e6 * (t2 - t0) / 2000:.1f} us per call until the GPU is done")
