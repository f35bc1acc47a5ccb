"""Python caller of the drop-in decoder step — the counterpart of the reference's
examples/python/testFlashDecoder.py, with the extension call live (the reference has it commented out, :124-126)
and the comparison asserted instead of printed.

    python examples/python/testFlashDecoder.py         (needs a B200; there is no CPU fallback)

The check is the caller's own naive attention in PyTorch on the GPU tensors: interleaved-pair RoPE on q and the new
k at position seq_len, append to the cache, softmax(q K^T / sqrt(d)) V over seq_len + 1 keys.
"""
import math
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import star_flash_attn  # noqa: E402  (built by `python starflashattention_b200/build.py --torch-ext`)


def rope_interleaved(x, pos, base=10000.0):
    """x (..., D): rotate pairs (x[2i], x[2i+1]) by pos / base^(2i/D)."""
    d = x.shape[-1]
    inv = base ** (-torch.arange(0, d, 2, dtype=torch.float32, device=x.device) / d)
    ang = float(pos) * inv
    c, s = ang.cos(), ang.sin()
    xf = x.float()
    even, odd = xf[..., 0::2], xf[..., 1::2]
    out = torch.stack((even * c - odd * s, odd * c + even * s), dim=-1).flatten(-2)
    return out.to(x.dtype)


def naive_step(qkv, k_cache, v_cache, seq_len, layer):
    B, _, H, D = qkv.shape
    outs = []
    for b in range(B):
        n = int(seq_len[b])
        q = rope_interleaved(qkv[b, 0], n)
        k_new = rope_interleaved(qkv[b, 1], n)
        k = torch.cat((k_cache[b, layer, :n], k_new[None]), 0).float()       # (n+1, H, D)
        v = torch.cat((v_cache[b, layer, :n], qkv[b, 2][None]), 0).float()
        s = torch.einsum("hd,nhd->hn", q.float(), k) / math.sqrt(D)
        outs.append(torch.einsum("hn,nhd->hd", s.softmax(-1), v))
    return torch.stack(outs)


def main():
    if not torch.cuda.is_available():
        raise SystemExit("no CUDA device: star_flash_attn has no CPU fallback")
    torch.manual_seed(0)
    batch_size, num_heads, head_dim = 1, 32, 128          # the reference example's shape (:97-105)
    memory_max_len, num_layer, idx_layer, cached = 1024, 2, 1, 511
    dev, dt = "cuda", torch.float16
    qkv = torch.randn(batch_size, 3, num_heads, head_dim, device=dev, dtype=dt)
    k_cache = torch.randn(batch_size, num_layer, memory_max_len, num_heads, head_dim, device=dev, dtype=dt)
    v_cache = torch.randn_like(k_cache)
    seq_len = torch.full((batch_size,), cached, dtype=torch.int32, device=dev)
    o = torch.empty(batch_size, num_heads, head_dim, device=dev, dtype=dt)
    bias = torch.empty(0, device=dev, dtype=dt)

    want = naive_step(qkv, k_cache, v_cache, seq_len.cpu(), idx_layer)
    got = star_flash_attn.mha_fwd_cuda(qkv, bias, bias, bias, k_cache, v_cache, seq_len, o, batch_size, memory_max_len,
                                       num_heads, head_dim, head_dim, memory_max_len, num_layer, idx_layer)
    torch.cuda.synchronize()
    err = (got.float() - want).abs().max().item()
    app = (k_cache[0, idx_layer, cached].float() - rope_interleaved(qkv[0, 1], cached).float()).abs().max().item()
    print(f"mha_fwd_cuda vs naive attention: max |o - o_ref| = {err:.3e}; appended k row error {app:.3e}")
    assert err <= 3e-3 and app <= 2e-3
    print("ok")


if __name__ == "__main__":
    main()
