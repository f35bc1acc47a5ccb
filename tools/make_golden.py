"""Generate tests/golden/*.npz from the reference's own Python code (run in the build container).

Imports ``LlamaAttention`` from /root/reference/examples/python/testFlashDecoder.py on the CPU
(the script imports the ``star_flash_attn`` extension at line 3, which is stubbed here; the
extension call itself is commented out in the reference, :124-126) and records its output on
small seeded inputs.  The reference cannot travel to the GPU box, the fixtures can.

    python tools/make_golden.py            # rewrites tests/golden/ref_decode_*.npz
"""
import contextlib
import io
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference/examples/python"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")


def load_reference():
    sys.modules.setdefault("star_flash_attn", types.ModuleType("star_flash_attn"))
    sys.path.insert(0, REF)
    import testFlashDecoder as ref  # noqa
    return ref


def run_ref(ref, qkv, kc, vc, cur_len, layer):
    attn = ref.LlamaAttention(qkv.shape[-1])
    with contextlib.redirect_stdout(io.StringIO()):      # the script prints whole tensors
        return attn.forward(qkv, kc, vc, cur_len, layer)


def main():
    ref = load_reference()
    os.makedirs(OUT, exist_ok=True)
    cases = [
        # name, B, H, D, L, M, current_seq_len (= cached tokens + 1), layer, dtype, fill
        ("ones_fp16", 1, 32, 128, 4, 1024, 512, 0, torch.float16, "ones"),     # the example's own shapes (:97-118)
        ("rand_fp32_d128", 2, 4, 128, 2, 96, 77, 1, torch.float32, "randn"),
        ("rand_fp32_d64", 1, 8, 64, 1, 160, 130, 0, torch.float32, "randn"),
        ("rand_fp16_d128", 2, 4, 128, 2, 300, 257, 1, torch.float16, "randn"),
        ("rand_bf16_d128", 1, 4, 128, 1, 64, 1, 0, torch.bfloat16, "randn"),   # empty cache: only the new token
    ]
    for i, (name, B, H, D, L, M, cur, layer, dt, fill) in enumerate(cases):
        g = torch.Generator().manual_seed(1000 + i)
        if fill == "ones":
            qkv = torch.ones(B, 3, H, D, dtype=dt)
            kc = torch.ones(B, L, M, H, D, dtype=dt)
            vc = torch.ones(B, L, M, H, D, dtype=dt)
        else:
            qkv = torch.randn(B, 3, H, D, generator=g).to(dt)
            kc = torch.randn(B, L, M, H, D, generator=g).to(dt)
            vc = torch.randn(B, L, M, H, D, generator=g).to(dt)
        run_dt = torch.float32 if dt != torch.float32 else dt
        # fp16/bf16 matmul on CPU follows the same formula; run the reference in fp32 on the
        # rounded inputs AND (for fp16) natively, keep both.
        o32 = run_ref(ref, qkv.to(run_dt), kc.to(run_dt), vc.to(run_dt), cur, layer)
        o_native = run_ref(ref, qkv, kc, vc, cur, layer).float() if dt != torch.float32 else o32
        if fill == "ones":                                  # keep the big all-ones fixture tiny on disk
            kc_s = np.zeros((0,), np.float32)
            vc_s = np.zeros((0,), np.float32)
            qkv_s = np.zeros((0,), np.float32)
        else:
            st = torch.float16 if dt == torch.float16 else torch.float32   # exact either way
            kc_s = kc[:, layer, :cur - 1].to(st).numpy()
            vc_s = vc[:, layer, :cur - 1].to(st).numpy()
            qkv_s = qkv.to(st).numpy()
        np.savez_compressed(
            os.path.join(OUT, f"ref_decode_{name}.npz"),
            qkv=qkv_s, k_cache_layer=kc_s, v_cache_layer=vc_s,
            meta=np.array([B, H, D, L, M, cur, layer], np.int64),
            dtype=str(dt).replace("torch.", ""), fill=fill,
            o_ref_fp32=o32.float().numpy(),            # (B,H,1,D): LlamaAttention output layout
            o_ref_native=o_native.numpy(),
        )
        print(name, tuple(o32.shape), float(o32.abs().max()))


if __name__ == "__main__":
    main()
