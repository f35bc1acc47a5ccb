"""CPU oracle for the attention forward / split-KV decode hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``starflashattention_b200/`` may import this
module; only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs do, and there only as the checker / the timed CPU baseline.

It restates, on the CPU, the naive ``softmax(Q K^T / sqrt(d)) V`` attention that the
reference repository's own Python example executes
(``/root/reference/examples/python/testFlashDecoder.py``):

* ``rope_cos_sin``            <- ``LlamaRotaryEmbedding``      (testFlashDecoder.py:7-26)
* ``rotate_pairs``            <- ``rotate_half``               (testFlashDecoder.py:28-43)
* ``apply_rope``              <- ``apply_rotary_pos_emb``      (testFlashDecoder.py:45-58)
* ``decode_step_reference``   <- ``LlamaAttention.forward``    (testFlashDecoder.py:61-94)
* ``attention_fwd``           <- lines :84 (scores), :88 (softmax), :92 (probs @ V), plus a
                                 bottom-right aligned causal mask and LSE = logsumexp(scores).

Parity pin: the reference holds no golden vectors or known-answer tests for this path
(SURVEY.md section 8c).  The oracle is therefore pinned against outputs of the reference's
own Python code imported in the build container (``tools/make_golden.py`` ->
``tests/golden/*.npz``) and against the one fixture the reference's examples imply
(all-ones inputs => O == 1.0, examples/cpp/testFlashDecoder.cc:63-78).  The reference's
CUDA kernel itself is not a usable numeric oracle.
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch

__all__ = [
    "rope_cos_sin",
    "rotate_pairs",
    "apply_rope",
    "attention_fwd",
    "decode_attention",
    "decode_step_reference",
    "combine_splits",
]


# --------------------------------------------------------------------------- RoPE
def rope_cos_sin(position: int, dim: int, base: float = 10000.0,
                 dtype: torch.dtype = torch.float32) -> Tuple[torch.Tensor, torch.Tensor]:
    """cos/sin of the interleaved-pair rotary angles at one position.

    angle_i = position / base^(2i/dim), repeated for both members of pair i
    (testFlashDecoder.py:10,20-25; kernel: src/flash_attn.cu:161-165,512-529).
    Returns two (dim,) tensors.
    """
    inv_freq = 1.0 / (base ** (torch.arange(0, dim, 2, dtype=torch.float32) / dim))
    ang = (torch.tensor(float(position), dtype=torch.float32) * inv_freq)
    ang = ang.repeat_interleave(2)
    return ang.cos().to(dtype), ang.sin().to(dtype)


def rotate_pairs(x: torch.Tensor) -> torch.Tensor:
    """(x0,x1,x2,x3,...) -> (-x1,x0,-x3,x2,...)   (testFlashDecoder.py:28-43)."""
    out = torch.empty_like(x)
    out[..., 0::2] = -x[..., 1::2]
    out[..., 1::2] = x[..., 0::2]
    return out


def apply_rope(x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor) -> torch.Tensor:
    """x*cos + rotate_pairs(x)*sin in fp32, cast back to x.dtype (testFlashDecoder.py:45-58)."""
    xf = x.float()
    return (xf * cos.float() + rotate_pairs(xf) * sin.float()).to(x.dtype)


# --------------------------------------------------------------------------- prefill
def attention_fwd(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, causal: bool = False,
                  softmax_scale: Optional[float] = None,
                  acc_dtype: torch.dtype = torch.float32,
                  chunk: int = 1024) -> Tuple[torch.Tensor, torch.Tensor]:
    """Naive attention on CPU.

    q: (B, Sq, Hq, D); k, v: (B, Sk, Hkv, D) with Hq % Hkv == 0 (GQA: q head h reads kv head
    h // (Hq/Hkv)).  Inputs are taken as they are (already rounded to fp16/bf16 by the caller)
    and up-cast to ``acc_dtype``; nothing is rounded in between.
    Causal mask is bottom-right aligned: query i sees keys j <= i + (Sk - Sq).
    Returns (O (B,Sq,Hq,D) acc_dtype, LSE (B,Hq,Sq) acc_dtype, natural log).
    Rows with no visible key give O = 0, LSE = -inf.
    Queries are processed in chunks so N=65536 slices do not materialise the full score matrix.
    """
    B, Sq, Hq, D = q.shape
    _, Sk, Hkv, _ = k.shape
    assert Hq % Hkv == 0
    g = Hq // Hkv
    scale = (1.0 / math.sqrt(D)) if softmax_scale is None else float(softmax_scale)
    qf = q.to(acc_dtype).permute(0, 2, 1, 3)                      # (B,Hq,Sq,D)
    kf = k.to(acc_dtype).permute(0, 2, 1, 3).repeat_interleave(g, dim=1)
    vf = v.to(acc_dtype).permute(0, 2, 1, 3).repeat_interleave(g, dim=1)
    out = torch.empty(B, Hq, Sq, D, dtype=acc_dtype)
    lse = torch.empty(B, Hq, Sq, dtype=acc_dtype)
    kt = kf.transpose(2, 3)
    key_idx = torch.arange(Sk)
    for s0 in range(0, Sq, chunk):
        s1 = min(Sq, s0 + chunk)
        scores = torch.matmul(qf[:, :, s0:s1], kt) * scale          # testFlashDecoder.py:84
        if causal:
            lim = torch.arange(s0, s1) + (Sk - Sq)
            mask = key_idx[None, :] > lim[:, None]
            scores = scores.masked_fill(mask, float("-inf"))
        l = torch.logsumexp(scores, dim=-1)
        probs = torch.softmax(scores, dim=-1)                       # testFlashDecoder.py:88
        probs = torch.nan_to_num(probs, nan=0.0)                    # fully masked rows
        out[:, :, s0:s1] = torch.matmul(probs, vf)                  # testFlashDecoder.py:92
        lse[:, :, s0:s1] = l
    return out.permute(0, 2, 1, 3).contiguous(), lse


# --------------------------------------------------------------------------- decode
def decode_attention(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                     seqlens: torch.Tensor, softmax_scale: Optional[float] = None,
                     acc_dtype: torch.dtype = torch.float32) -> Tuple[torch.Tensor, torch.Tensor]:
    """Single-token decode: q (B,Hq,D), caches (B,M,Hkv,D), seqlens (B,) = visible keys.

    Returns (O (B,Hq,D), LSE (B,Hq)) in ``acc_dtype``.  seqlens[b] == 0 gives O=0, LSE=-inf.
    """
    B, Hq, D = q.shape
    outs, lses = [], []
    for b in range(B):
        n = int(seqlens[b])
        o, l = attention_fwd(q[b:b + 1, None], k_cache[b:b + 1, :n], v_cache[b:b + 1, :n],
                             causal=False, softmax_scale=softmax_scale, acc_dtype=acc_dtype)
        outs.append(o[0, 0])
        lses.append(l[0, :, 0])
    return torch.stack(outs), torch.stack(lses)


def gather_pages(pool: torch.Tensor, block_table: torch.Tensor) -> torch.Tensor:
    """Paged cache -> the dense (B, blocks * page_size, Hkv, D) cache it stands for: pool (pages, page_size, Hkv, D),
    block_table (B, blocks) of pool pages.  The reference has no paged cache (its layout is the dense one of
    src/params.h:22-25); this is the definition the paged kernel is checked against."""
    B, blocks = block_table.shape
    _, ps, Hkv, D = pool.shape
    return pool[block_table.long().reshape(-1)].reshape(B, blocks * ps, Hkv, D)


def decode_step_reference(qkv: torch.Tensor, k_cache_table: torch.Tensor,
                          v_cache_table: torch.Tensor, seq_len: torch.Tensor, idx_layer: int,
                          rotary_dim: Optional[int] = None,
                          table_dtype: Optional[torch.dtype] = None,
                          acc_dtype: torch.dtype = torch.float32):
    """The reference decoder step (intended algorithm).

    qkv (B,3,H,D); caches (B,L,M,H,D); seq_len (B,) int = tokens already cached.  For each
    batch: rotate q,k at position seq_len[b] (src/flash_attn.cu:600-601,685-687), append k,v at
    cache row seq_len[b] (:696-701), attend over seq_len[b]+1 keys (:571), scale 1/sqrt(D)
    (src/flash_api.cpp:27).  Mirrors LlamaAttention.forward(…, current_seq_len=seq_len+1, …)
    (testFlashDecoder.py:61-94) where RoPE is computed in fp32 and cast back to the input dtype
    before the matmuls.  ``table_dtype`` rounds cos/sin first (the fp16 tables that
    compute_rotary_table<half> produces).
    Returns (O (B,H,D) acc_dtype, LSE (B,H), k_new (B,H,D), v_new (B,H,D)) and does NOT mutate
    the caches.
    """
    B, three, H, D = qkv.shape
    assert three == 3
    rd = D if rotary_dim is None else rotary_dim
    outs, lses, kn, vn = [], [], [], []
    for b in range(B):
        pos = int(seq_len[b])
        q, k, v = qkv[b, 0], qkv[b, 1], qkv[b, 2]               # (H,D)
        cos, sin = rope_cos_sin(pos, rd)
        if table_dtype is not None:
            cos, sin = cos.to(table_dtype), sin.to(table_dtype)
        q = q.clone()
        k = k.clone()
        q[:, :rd] = apply_rope(q[:, :rd], cos, sin)
        k[:, :rd] = apply_rope(k[:, :rd], cos, sin)
        keys = torch.cat([k_cache_table[b, idx_layer, :pos], k[None]], dim=0)   # (pos+1,H,D)
        vals = torch.cat([v_cache_table[b, idx_layer, :pos], v[None]], dim=0)
        o, l = attention_fwd(q[None, None], keys[None], vals[None], acc_dtype=acc_dtype)
        outs.append(o[0, 0])
        lses.append(l[0, :, 0])
        kn.append(k)
        vn.append(v)
    return torch.stack(outs), torch.stack(lses), torch.stack(kn), torch.stack(vn)


def decode_step_gqa(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                    seq_len: torch.Tensor, rotary_dim: Optional[int] = None, table_dtype: Optional[torch.dtype] = None,
                    softmax_scale: Optional[float] = None):
    """The decoder step of decode_step_reference for grouped query heads and one dense cache per call:
    q (B,Hq,D), k / v (B,Hkv,D) of the new token, caches (B,M,Hkv,D), seq_len (B,) tokens already cached.
    Same order of operations (RoPE at position seq_len[b] in fp32 and cast back, append at row seq_len[b],
    attention over seq_len[b] + 1 keys; src/flash_attn.cu:591-601,696-701, testFlashDecoder.py:61-94).
    Returns (O (B,Hq,D), LSE (B,Hq), rotated k (B,Hkv,D)); the caches are not mutated."""
    B, Hq, D = q.shape
    rd = D if rotary_dim is None else rotary_dim
    outs, lses, kn = [], [], []
    for b in range(B):
        pos = int(seq_len[b])
        cos, sin = rope_cos_sin(pos, rd)
        if table_dtype is not None:
            cos, sin = cos.to(table_dtype), sin.to(table_dtype)
        qb, kb = q[b].clone(), k[b].clone()
        qb[:, :rd] = apply_rope(qb[:, :rd], cos, sin)
        kb[:, :rd] = apply_rope(kb[:, :rd], cos, sin)
        keys = torch.cat([k_cache[b, :pos], kb[None]], dim=0)
        vals = torch.cat([v_cache[b, :pos], v[b][None]], dim=0)
        o, l = attention_fwd(qb[None, None], keys[None], vals[None], softmax_scale=softmax_scale)
        outs.append(o[0, 0])
        lses.append(l[0, :, 0])
        kn.append(kb)
    return torch.stack(outs), torch.stack(lses), torch.stack(kn)


def combine_splits(o_part: torch.Tensor, m_part: torch.Tensor, l_part: torch.Tensor):
    """Merge split-KV partials (intended form of flash_combine_kernel, src/flash_attn.cu:877-935).

    o_part (..., S, D) un-normalised partial outputs, m_part (..., S) running maxima (natural-log
    domain), l_part (..., S) partial denominators.  M = max m_s; L = sum e^{m_s-M} l_s;
    O = sum e^{m_s-M} O_s / L; LSE = M + log L.
    """
    M = m_part.max(dim=-1, keepdim=True).values
    M0 = torch.where(torch.isinf(M), torch.zeros_like(M), M)
    w = torch.exp(m_part - M0)
    w = torch.where(torch.isinf(m_part), torch.zeros_like(w), w)
    L = (w * l_part).sum(-1)
    O = (w[..., None] * o_part).sum(-2)
    O = torch.where(L[..., None] > 0, O / L[..., None].clamp_min(1e-300), torch.zeros_like(O))
    lse = torch.where(L > 0, M0[..., 0] + torch.log(L.clamp_min(1e-300)),
                      torch.full_like(L, float("-inf")))
    return O, lse
