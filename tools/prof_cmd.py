"""Short command lines for ncu (a few launches each): python tools/prof_cmd.py fwd|fwd64|decode|step"""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import starflashattention_b200 as sfa  # noqa: E402

what = sys.argv[1] if len(sys.argv) > 1 else "fwd"
g = torch.Generator(device="cuda").manual_seed(0)
rnd = lambda shape, dt: torch.randn(shape, generator=g, device="cuda", dtype=torch.float32).to(dt)
n = 4
if what == "fwd":        # C3
    q, k, v = (rnd((8, 8192, 32, 128), torch.bfloat16) for _ in range(3))
    for _ in range(n):
        sfa.fwd(q, k, v, causal=True)
elif what == "fwd64":    # C2
    q, k, v = (rnd((4, 4096, 16, 64), torch.float16) for _ in range(3))
    for _ in range(n):
        sfa.fwd(q, k, v, causal=False)
elif what == "decode":   # C4
    q, k, v = rnd((8, 32, 128), torch.bfloat16), rnd((8, 32768, 8, 128), torch.bfloat16), rnd((8, 32768, 8, 128), torch.bfloat16)
    sl = torch.full((8,), 32768, dtype=torch.int32, device="cuda")
    for _ in range(n):
        sfa.decode(q, k, v, sl)
else:                    # drop-in decoder step, fused RoPE + append
    B, L, M, H, D = 8, 1, 32768, 8, 128
    kc, vc = rnd((B, L, M, H, D), torch.bfloat16), rnd((B, L, M, H, D), torch.bfloat16)
    qkv = rnd((B, 3, H, D), torch.bfloat16)
    cos, sin = sfa.compute_rotary_table(M, D, dtype=torch.bfloat16)
    seq = torch.full((B,), M - 64, dtype=torch.int32, device="cuda")
    o = torch.empty((B, H, D), dtype=torch.bfloat16, device="cuda")
    for _ in range(n):
        sfa.mha_fwd_cuda(qkv, None, None, None, kc, vc, seq, o, B, M, H, D, D, M, L, 0, rotary_cos_table=cos, rotary_sin_table=sin)
torch.cuda.synchronize()
print("ok", what, sfa.launch_count())
