"""Host-side mirror of the reference's operator interface, on top of the C ABI.

* ``mha_fwd_cuda``          <- ``star_flash_attn.mha_fwd_cuda`` (reference src/flash_api.cpp:42-68, bound at :74-79):
                               same positional arguments, ``o`` written in place and returned, biases accepted
                               and ignored, work queued on the current CUDA stream.
* ``compute_rotary_table``  <- ``compute_rotary_table<half>`` (reference src/flash_attn.cu:531-538)
* ``init_half_array``       <- ``init_half_array`` (reference src/flash_attn.cu:502-510)
* ``fwd`` / ``decode``      additions: prefill forward and GQA/bf16 split-KV decode with LSE (no reference
                               counterpart; arithmetic of examples/python/testFlashDecoder.py:84-92).

PyTorch is used for device memory and streams only; every computation happens in libStarFlashAttention.so.
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch

from . import _capi
from ._capi import DecodeArgs, DecodePagedArgs, DecoderStepArgs, DecoderStepPagedArgs, FwdArgs, check, lib

__all__ = ["DecoderStepGraph", "fwd", "decode", "fwd_host", "decode_host", "mha_fwd_cuda", "compute_rotary_table", "init_half_array", "fwd_watchdog",
           "launch_count", "decode_num_splits", "tuning_set"]


def _dtype_code(t: torch.Tensor) -> int:
    if t.dtype == torch.float16:
        return _capi.SFA_FP16
    if t.dtype == torch.bfloat16:
        return _capi.SFA_BF16
    raise TypeError(f"only float16 and bfloat16 are supported, got {t.dtype}")


def _require_cuda(*tensors: torch.Tensor) -> torch.device:
    dev = None
    for t in tensors:
        if t is None:
            continue
        if not t.is_cuda:
            raise ValueError("all tensors must live on a CUDA device (there is no CPU fallback)")
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise ValueError("all tensors must be on the same device")
    return dev


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def fwd(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, causal: bool = False,
        softmax_scale: Optional[float] = None, out: Optional[torch.Tensor] = None,
        return_lse: bool = True) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """O = softmax(Q K^T * scale [+ causal]) V.  q (B,Sq,Hq,D); k, v (B,Sk,Hkv,D); last dim contiguous.

    Returns (out like q, lse (B,Hq,Sq) float32 or None).  Causal mask is bottom-right aligned.
    """
    dev = _require_cuda(q, k, v, out)
    if q.dim() != 4 or k.dim() != 4 or v.dim() != 4:
        raise ValueError("q, k, v must be 4-D (batch, seqlen, heads, head_dim)")
    B, Sq, Hq, D = q.shape
    Bk, Sk, Hkv, Dk = k.shape
    if v.shape != k.shape or Bk != B or Dk != D:
        raise ValueError("k and v must have equal shapes matching q's batch and head_dim")
    if q.dtype != k.dtype or q.dtype != v.dtype:
        raise TypeError("q, k, v must share one dtype")
    for t in (q, k, v):
        if t.stride(-1) != 1:
            raise ValueError("the head_dim axis must be contiguous")
    if out is None:
        out = torch.empty((B, Sq, Hq, D), dtype=q.dtype, device=dev)
    elif out.shape != q.shape or out.dtype != q.dtype or out.stride(-1) != 1:
        raise ValueError("out must look like q")
    lse = torch.empty((B, Hq, Sq), dtype=torch.float32, device=dev) if return_lse else None
    a = FwdArgs()
    a.q, a.k, a.v, a.o = q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr()
    a.lse = lse.data_ptr() if lse is not None else None
    a.batch, a.seqlen_q, a.seqlen_k, a.heads_q, a.heads_kv, a.head_dim = B, Sq, Sk, Hq, Hkv, D
    a.q_batch_stride, a.q_row_stride, a.q_head_stride = q.stride(0), q.stride(1), q.stride(2)
    a.k_batch_stride, a.k_row_stride, a.k_head_stride = k.stride(0), k.stride(1), k.stride(2)
    a.v_batch_stride, a.v_row_stride, a.v_head_stride = v.stride(0), v.stride(1), v.stride(2)
    a.o_batch_stride, a.o_row_stride, a.o_head_stride = out.stride(0), out.stride(1), out.stride(2)
    a.softmax_scale = float(softmax_scale) if softmax_scale is not None else 0.0
    a.causal = 1 if causal else 0
    a.dtype = _dtype_code(q)
    with torch.cuda.device(dev):
        check(lib.sfa_fwd(a, _stream()))
    return out, lse


def _decode_args(q, k_cache, v_cache, seqlens, out, lse, softmax_scale, num_splits) -> DecodeArgs:
    B, Hq, D = q.shape
    _, M, Hkv, _ = k_cache.shape
    a = DecodeArgs()
    a.q, a.k_cache, a.v_cache = q.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr()
    a.seqlens = seqlens.data_ptr() if seqlens is not None else None
    a.o = out.data_ptr()
    a.lse = lse.data_ptr() if lse is not None else None
    a.batch, a.seqlen_k, a.heads_q, a.heads_kv, a.head_dim = B, M, Hq, Hkv, D
    a.q_batch_stride, a.q_head_stride = q.stride(0), q.stride(1)
    a.k_batch_stride, a.k_row_stride, a.k_head_stride = k_cache.stride(0), k_cache.stride(1), k_cache.stride(2)
    a.v_batch_stride, a.v_row_stride, a.v_head_stride = v_cache.stride(0), v_cache.stride(1), v_cache.stride(2)
    a.o_batch_stride, a.o_head_stride = out.stride(0), out.stride(1)
    a.softmax_scale = float(softmax_scale) if softmax_scale is not None else 0.0
    a.num_splits = int(num_splits)
    a.dtype = _dtype_code(q)
    a.workspace, a.workspace_bytes = None, 0
    return a


def decode(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
           seqlens: Optional[torch.Tensor] = None, softmax_scale: Optional[float] = None,
           num_splits: int = 0, out: Optional[torch.Tensor] = None,
           return_lse: bool = True) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """Single-token attention over a KV cache.  q (B,Hq,D); caches (B,M,Hkv,D); seqlens int32 (B) or None."""
    dev = _require_cuda(q, k_cache, v_cache, seqlens, out)
    if q.dim() != 3 or k_cache.dim() != 4 or v_cache.shape != k_cache.shape:
        raise ValueError("q must be (batch, heads, head_dim); caches (batch, max_len, heads_kv, head_dim)")
    if q.shape[0] != k_cache.shape[0] or q.shape[2] != k_cache.shape[3]:
        raise ValueError("batch / head_dim mismatch between q and the caches")
    if q.dtype != k_cache.dtype or q.dtype != v_cache.dtype:
        raise TypeError("q and the caches must share one dtype")
    for t in (q, k_cache, v_cache):
        if t.stride(-1) != 1:
            raise ValueError("the head_dim axis must be contiguous")
    if seqlens is not None and (seqlens.dtype != torch.int32 or seqlens.numel() != q.shape[0] or not seqlens.is_contiguous()):
        raise ValueError("seqlens must be a contiguous int32 tensor of shape (batch,)")
    if out is None:
        out = torch.empty_like(q, memory_format=torch.contiguous_format)
    elif out.shape != q.shape or out.dtype != q.dtype or out.stride(-1) != 1:
        raise ValueError("out must look like q")
    lse = torch.empty(q.shape[:2], dtype=torch.float32, device=dev) if return_lse else None
    a = _decode_args(q, k_cache, v_cache, seqlens, out, lse, softmax_scale, num_splits)
    with torch.cuda.device(dev):
        check(lib.sfa_decode(a, _stream()))
    return out, lse


def decode_paged(q: torch.Tensor, k_pool: torch.Tensor, v_pool: torch.Tensor, block_table: torch.Tensor,
                 seqlens: torch.Tensor, softmax_scale: Optional[float] = None, num_splits: int = 0,
                 out: Optional[torch.Tensor] = None,
                 return_lse: bool = True) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """Single-token attention over a PAGED KV cache (scope row f1).  q (B,Hq,D); pools (pages, page_size, Hkv, D);
    block_table int32 (B, blocks): pool page of each page_size-key block of a sequence; seqlens int32 (B) visible
    keys, at most blocks * page_size.  Same kernel and arithmetic as :func:`decode`."""
    dev = _require_cuda(q, k_pool, v_pool, block_table, seqlens, out)
    if q.dim() != 3 or k_pool.dim() != 4 or v_pool.shape != k_pool.shape:
        raise ValueError("q must be (batch, heads, head_dim); pools (pages, page_size, heads_kv, head_dim)")
    if q.shape[2] != k_pool.shape[3]:
        raise ValueError("head_dim mismatch between q and the pools")
    if q.dtype != k_pool.dtype or q.dtype != v_pool.dtype:
        raise TypeError("q and the pools must share one dtype")
    for t in (q, k_pool, v_pool):
        if t.stride(-1) != 1:
            raise ValueError("the head_dim axis must be contiguous")
    if block_table.dtype != torch.int32 or block_table.dim() != 2 or block_table.shape[0] != q.shape[0] or \
            block_table.stride(1) != 1:
        raise ValueError("block_table must be int32 (batch, blocks) with contiguous rows")
    if seqlens is None or seqlens.dtype != torch.int32 or seqlens.numel() != q.shape[0] or not seqlens.is_contiguous():
        raise ValueError("seqlens must be a contiguous int32 tensor of shape (batch,)")
    if out is None:
        out = torch.empty_like(q, memory_format=torch.contiguous_format)
    elif out.shape != q.shape or out.dtype != q.dtype or out.stride(-1) != 1:
        raise ValueError("out must look like q")
    lse = torch.empty(q.shape[:2], dtype=torch.float32, device=dev) if return_lse else None
    pa = DecodePagedArgs()
    pa.base = _decode_args(q, k_pool, v_pool, seqlens, out, lse, softmax_scale, num_splits)
    pa.base.seqlen_k = block_table.shape[1] * k_pool.shape[1]      # largest addressable length
    pa.block_table, pa.block_table_stride = block_table.data_ptr(), block_table.stride(0)
    pa.page_size, pa.num_pages = k_pool.shape[1], k_pool.shape[0]
    with torch.cuda.device(dev):
        check(lib.sfa_decode_paged(pa, _stream()))
    return out, lse


def decode_num_splits(batch: int, seqlen_k: int, heads_kv: int, device=None) -> int:
    """Number of KV splits the launcher picks for this shape on the current device."""
    a = DecodeArgs()
    a.batch, a.seqlen_k, a.heads_kv, a.num_splits = batch, seqlen_k, heads_kv, 0
    with torch.cuda.device(device if device is not None else torch.cuda.current_device()):
        n = lib.sfa_decode_num_splits(a)
    if n < 0:
        check(n)
    return n


def mha_fwd_cuda(qkv: torch.Tensor, q_bias, k_bias, v_bias, k_cache_table: torch.Tensor,
                 v_cache_table: torch.Tensor, seq_len: torch.Tensor, o: torch.Tensor, batch_size: int,
                 memory_max_len: int, num_heads: int, head_dim: int, rotary_embedding_dim: int,
                 max_input_length: int, num_layer: int, idx_layer: int, *,
                 rotary_cos_table: Optional[torch.Tensor] = None, rotary_sin_table: Optional[torch.Tensor] = None,
                 lse: Optional[torch.Tensor] = None, num_splits: int = 0,
                 workspace: Optional[torch.Tensor] = None) -> torch.Tensor:
    """One decoder step: RoPE(q, k) at position seq_len[b], append k, v to layer ``idx_layer`` of the caches,
    attend q over the seq_len[b]+1 cached keys.  Mirrors the reference entry argument for argument
    (src/flash_api.cpp:42-58); the keyword-only arguments are additions (``workspace``: caller-owned uint8 device
    scratch of at least ``decoder_step_workspace_bytes`` bytes, instead of the library's per-stream scratch)."""
    dev = _require_cuda(qkv, k_cache_table, v_cache_table, seq_len, o, rotary_cos_table, rotary_sin_table, lse)
    if seq_len.dtype != torch.int32 or not seq_len.is_contiguous() or seq_len.numel() != batch_size:
        raise ValueError("seq_len must be a contiguous int32 tensor of shape (batch_size,)")
    want_cache = (batch_size, num_layer, memory_max_len, num_heads, head_dim)
    if tuple(k_cache_table.shape) != want_cache or tuple(v_cache_table.shape) != want_cache:
        raise ValueError(f"caches must have shape {want_cache}")
    if not (k_cache_table.is_contiguous() and v_cache_table.is_contiguous() and o.is_contiguous()):
        raise ValueError("caches and o must be contiguous")
    if tuple(qkv.shape) != (batch_size, 3, num_heads, head_dim) or qkv.stride(-1) != 1 or \
            qkv.stride(2) != head_dim or qkv.stride(1) != num_heads * head_dim:
        raise ValueError("qkv must be (batch_size, 3, num_heads, head_dim) with a dense token row")
    if tuple(o.shape) != (batch_size, num_heads, head_dim):
        raise ValueError("o must be (batch_size, num_heads, head_dim)")
    if not (qkv.dtype == k_cache_table.dtype == v_cache_table.dtype == o.dtype):
        raise TypeError("qkv, caches and o must share one dtype")
    for t in (rotary_cos_table, rotary_sin_table):
        if t is not None and (t.dtype != qkv.dtype or not t.is_contiguous() or
                              tuple(t.shape) != (memory_max_len, rotary_embedding_dim // 2)):
            raise ValueError("rotary tables must be contiguous (memory_max_len, rotary_embedding_dim/2) in the data dtype")
    a = DecoderStepArgs()
    a.qkv, a.o, a.seq_len = qkv.data_ptr(), o.data_ptr(), seq_len.data_ptr()
    a.k_cache_table, a.v_cache_table = k_cache_table.data_ptr(), v_cache_table.data_ptr()
    a.rotary_cos_table = rotary_cos_table.data_ptr() if rotary_cos_table is not None else None
    a.rotary_sin_table = rotary_sin_table.data_ptr() if rotary_sin_table is not None else None
    a.lse = lse.data_ptr() if lse is not None else None
    a.batch_size, a.memory_max_len, a.num_heads, a.head_dim = batch_size, memory_max_len, num_heads, head_dim
    a.head_dim_inv = 1.0 / math.sqrt(head_dim)                      # src/flash_api.cpp:27
    a.rotary_embedding_dim = rotary_embedding_dim
    a.qkv_stride = qkv.stride(0)
    a.num_layer, a.idx_layer = num_layer, idx_layer
    a.num_splits = int(num_splits)
    a.dtype = _dtype_code(qkv)
    if workspace is not None:
        a.workspace, a.workspace_bytes = workspace.data_ptr(), workspace.numel() * workspace.element_size()
    with torch.cuda.device(dev):
        check(lib.sfa_decoder_step(a, _stream()))
    return o


def decoder_step_workspace_bytes(batch_size: int, num_heads: int, head_dim: int) -> int:
    return int(lib.sfa_decoder_step_workspace_bytes(batch_size, num_heads, head_dim))


def decoder_step_paged(qkv: torch.Tensor, k_pool: torch.Tensor, v_pool: torch.Tensor, block_table: torch.Tensor,
                       seq_len: torch.Tensor, heads_q: int, o: Optional[torch.Tensor] = None, *,
                       rotary_embedding_dim: Optional[int] = None, rotary_cos_table: Optional[torch.Tensor] = None,
                       rotary_sin_table: Optional[torch.Tensor] = None, lse: Optional[torch.Tensor] = None,
                       softmax_scale: Optional[float] = None, num_splits: int = 0) -> torch.Tensor:
    """The decoder step of :func:`mha_fwd_cuda` over a PAGED cache, with GQA: RoPE(q, k) at position seq_len[b],
    append k, v to row seq_len[b] of the sequence (page ``block_table[b, seq_len[b] // page_size]``, assigned by the
    caller), attend over seq_len[b] + 1 keys - one fused kernel plus the split merge.
    qkv (B, heads_q + 2 heads_kv, D): q heads, k heads, v heads of the new token; pools (pages, page_size, Hkv, D);
    block_table int32 (B, blocks); seq_len int32 (B).  Returns o (B, heads_q, D)."""
    dev = _require_cuda(qkv, k_pool, v_pool, block_table, seq_len, o, rotary_cos_table, rotary_sin_table, lse)
    if k_pool.dim() != 4 or v_pool.shape != k_pool.shape or k_pool.stride() != v_pool.stride() or k_pool.stride(-1) != 1:
        raise ValueError("pools must be (pages, page_size, heads_kv, head_dim) with equal strides and a contiguous head_dim")
    pages, page, Hkv, D = k_pool.shape
    B = qkv.shape[0]
    if tuple(qkv.shape) != (B, heads_q + 2 * Hkv, D) or qkv.stride(-1) != 1 or qkv.stride(1) != D:
        raise ValueError("qkv must be (batch, heads_q + 2 * heads_kv, head_dim) with a dense token row")
    if not (qkv.dtype == k_pool.dtype == v_pool.dtype):
        raise TypeError("qkv and the pools must share one dtype")
    if block_table.dtype != torch.int32 or block_table.dim() != 2 or block_table.shape[0] != B or block_table.stride(1) != 1:
        raise ValueError("block_table must be int32 (batch, blocks) with contiguous rows")
    if seq_len.dtype != torch.int32 or not seq_len.is_contiguous() or seq_len.numel() != B:
        raise ValueError("seq_len must be a contiguous int32 tensor of shape (batch,)")
    if o is None:
        o = torch.empty((B, heads_q, D), dtype=qkv.dtype, device=dev)
    elif tuple(o.shape) != (B, heads_q, D) or o.dtype != qkv.dtype or not o.is_contiguous():
        raise ValueError("o must be a contiguous (batch, heads_q, head_dim) tensor of the data dtype")
    max_len = block_table.shape[1] * page
    rd = D if rotary_embedding_dim is None else rotary_embedding_dim
    for t in (rotary_cos_table, rotary_sin_table):
        if t is not None and (t.dtype != qkv.dtype or not t.is_contiguous() or t.dim() != 2 or
                              t.shape[0] < max_len or t.shape[1] != rd // 2):
            raise ValueError("rotary tables must be contiguous (>= blocks * page_size, rotary_embedding_dim/2) in the data dtype")
    a = DecoderStepPagedArgs()
    a.qkv, a.o, a.seq_len = qkv.data_ptr(), o.data_ptr(), seq_len.data_ptr()
    a.lse = lse.data_ptr() if lse is not None else None
    a.k_pool, a.v_pool, a.block_table = k_pool.data_ptr(), v_pool.data_ptr(), block_table.data_ptr()
    a.rotary_cos_table = rotary_cos_table.data_ptr() if rotary_cos_table is not None else None
    a.rotary_sin_table = rotary_sin_table.data_ptr() if rotary_sin_table is not None else None
    a.page_stride, a.row_stride, a.head_stride = k_pool.stride(0), k_pool.stride(1), k_pool.stride(2)
    a.block_table_stride, a.qkv_stride = block_table.stride(0), qkv.stride(0)
    a.batch_size, a.max_len, a.heads_q, a.heads_kv, a.head_dim = B, max_len, heads_q, Hkv, D
    a.page_size, a.num_pages = page, pages
    a.rotary_embedding_dim = rd
    a.softmax_scale = float(softmax_scale) if softmax_scale is not None else 0.0
    a.num_splits = int(num_splits)
    a.dtype = _dtype_code(qkv)
    with torch.cuda.device(dev):
        check(lib.sfa_decoder_step_paged(a, _stream()))
    return o


class DecoderStepGraph:
    """A whole decoder step (every layer's fused RoPE + append + split-KV attention + combine) captured once in a
    CUDA graph and replayed per token: the reference pays 7 cudaMallocs, 3 blocking copies and a device
    synchronise per call (src/flash_attn.cu:947-991); here a token costs one graph launch.

    The graph reads ``qkv`` (num_layer, batch, 3, heads, head_dim) and ``seq_len`` (batch) and writes ``o``
    (num_layer, batch, heads, head_dim); these are the instance's own static tensors - copy a token's
    projections into ``qkv`` and call :meth:`step`, which also advances ``seq_len`` on the device.
    (In a real model layer l's qkv depends on layer l-1's output; the projections between the attention calls
    would be captured in the same graph by the caller.  This class covers the attention path this library owns.)"""

    def __init__(self, k_cache_table: torch.Tensor, v_cache_table: torch.Tensor, seq_len: torch.Tensor, *,
                 rotary_embedding_dim: Optional[int] = None, rotary_cos_table: Optional[torch.Tensor] = None,
                 rotary_sin_table: Optional[torch.Tensor] = None, num_splits: int = 0):
        B, L, M, H, D = k_cache_table.shape
        dev = k_cache_table.device
        self.k_cache_table, self.v_cache_table = k_cache_table, v_cache_table
        self.seq_len = seq_len.to(device=dev, dtype=torch.int32).clone()
        self.qkv = torch.zeros((L, B, 3, H, D), dtype=k_cache_table.dtype, device=dev)
        self.o = torch.empty((L, B, H, D), dtype=k_cache_table.dtype, device=dev)
        self.lse = torch.empty((L, B, H), dtype=torch.float32, device=dev)
        rot = D if rotary_embedding_dim is None else rotary_embedding_dim
        # The graph owns its scratch: the captured launches keep its address for as long as the graph lives, whatever
        # the library's own per-stream scratch does in the meantime.
        self.workspace = torch.empty(decoder_step_workspace_bytes(B, H, D), dtype=torch.uint8, device=dev)

        def run():
            for layer in range(L):
                mha_fwd_cuda(self.qkv[layer], None, None, None, k_cache_table, v_cache_table, self.seq_len,
                             self.o[layer], B, M, H, D, rot, M, L, layer, rotary_cos_table=rotary_cos_table,
                             rotary_sin_table=rotary_sin_table, lse=self.lse[layer], num_splits=num_splits,
                             workspace=self.workspace)
            self.seq_len.add_(1)

        # Warm up on a side stream, then roll the state back.
        k0, v0, s0 = k_cache_table.clone(), v_cache_table.clone(), self.seq_len.clone()
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        n0 = launch_count()
        with torch.cuda.stream(side):
            run()
        self.launches_per_step = launch_count() - n0        # kernels of this library in one step
        torch.cuda.current_stream(dev).wait_stream(side)
        k_cache_table.copy_(k0); v_cache_table.copy_(v0); self.seq_len.copy_(s0)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            run()
        k_cache_table.copy_(k0); v_cache_table.copy_(v0); self.seq_len.copy_(s0)

    def step(self, qkv: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Replays the captured step on the current stream; returns the static output tensor."""
        if qkv is not None:
            self.qkv.copy_(qkv, non_blocking=True)
        self.graph.replay()
        return self.o


def compute_rotary_table(max_seq_len: int, rot_embed_dim: int, dtype=torch.float16, device="cuda"):
    """cos / sin tables of shape (max_seq_len, rot_embed_dim // 2): angle = pos / 10000^(2i/rot_embed_dim)."""
    cos = torch.empty((max_seq_len, rot_embed_dim // 2), dtype=dtype, device=device)
    sin = torch.empty_like(cos)
    with torch.cuda.device(cos.device):
        check(lib.sfa_rotary_table(cos.data_ptr(), sin.data_ptr(), max_seq_len, rot_embed_dim, _dtype_code(cos), _stream()))
    return cos, sin


def init_half_array(array: torch.Tensor, value: float) -> torch.Tensor:
    """array[:] = value (the reference's test utility); fp16 or bf16, contiguous."""
    _require_cuda(array)
    if not array.is_contiguous():
        raise ValueError("array must be contiguous")
    with torch.cuda.device(array.device):
        check(lib.sfa_fill(array.data_ptr(), float(value), array.numel(), _dtype_code(array), _stream()))
    return array


def fwd_watchdog(clear: bool = True) -> int:
    """Synchronises and returns the prefill kernel's pipeline watchdog code (0 = healthy)."""
    rc = lib.sfa_fwd_watchdog(1 if clear else 0)
    if rc < 0:
        check(rc)
    return rc


def launch_count() -> int:
    return int(lib.sfa_launch_count())


def tuning_set(name: str, value: int) -> None:
    """Test / tuning aid: one of the library's A-B switches (see sfa_tuning_set in include/star_flash_attn.h)."""
    check(lib.sfa_tuning_set(name.encode(), int(value)))


# ------------------------------------------------------------------------------------------------
# Host-buffer entry points: pinned host tensors in, pinned host tensors out.  The pipeline itself (piece plan, three
# streams, device staging) lives in the library (sfa_fwd_host / sfa_decode_host, csrc/host_pipeline.cu); these are
# thin callers, as the reference's Python layer is over its C++ host layer (src/flash_api.cpp:42-68).
def _pinned(t: torch.Tensor, what: str):
    if t.is_cuda:
        raise ValueError(f"{what} must be a host tensor (use fwd()/decode() for device tensors)")
    if not t.is_pinned():
        raise ValueError(f"{what} must live in pinned host memory (tensor.pin_memory())")


def fwd_host(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, causal: bool = False,
             softmax_scale: Optional[float] = None, out: Optional[torch.Tensor] = None,
             lse: Optional[torch.Tensor] = None, device=None, chunks: Optional[int] = None):
    """Prefill forward on pinned HOST tensors: q (B,Sq,Hq,D), k/v (B,Sk,Hkv,D) -> (out, lse) in pinned host memory.
    q/k/v/out may be slices of the heads of larger tensors (a rank's share); lse is dense (B,Hq,Sq).
    ``chunks``: number of uniform pieces the call is cut into (default: the library's own plan, see sfa_fwd_host in
    include/star_flash_attn.h).
    Asynchronous: the results are complete once the current stream of ``device`` has been synchronised."""
    for t, n in ((q, "q"), (k, "k"), (v, "v")):
        _pinned(t, n)
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    B, Sq, Hq, D = q.shape
    if out is None:
        out = torch.empty(q.shape, dtype=q.dtype, pin_memory=True)
    if lse is None:
        lse = torch.empty((B, Hq, Sq), dtype=torch.float32, pin_memory=True)
    _pinned(out, "out"); _pinned(lse, "lse")
    if not lse.is_contiguous() or tuple(lse.shape) != (B, Hq, Sq) or lse.dtype != torch.float32:
        raise ValueError("lse must be a contiguous float32 (batch, heads_q, seqlen_q) tensor")
    if q.dtype != k.dtype or q.dtype != v.dtype or out.dtype != q.dtype or out.shape != q.shape:
        raise TypeError("q, k, v, out must share one dtype and out must look like q")
    for t in (q, k, v, out):
        if t.stride(3) != 1:
            raise ValueError("the head_dim axis must be contiguous")
    h = _capi.FwdHostArgs()
    a = h.t
    a.q, a.k, a.v, a.o, a.lse = q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(), lse.data_ptr()
    a.batch, a.seqlen_q, a.seqlen_k, a.heads_q, a.heads_kv, a.head_dim = B, Sq, k.shape[1], Hq, k.shape[2], D
    a.q_batch_stride, a.q_row_stride, a.q_head_stride = q.stride(0), q.stride(1), q.stride(2)
    a.k_batch_stride, a.k_row_stride, a.k_head_stride = k.stride(0), k.stride(1), k.stride(2)
    a.v_batch_stride, a.v_row_stride, a.v_head_stride = v.stride(0), v.stride(1), v.stride(2)
    a.o_batch_stride, a.o_row_stride, a.o_head_stride = out.stride(0), out.stride(1), out.stride(2)
    a.softmax_scale = float(softmax_scale) if softmax_scale is not None else 0.0
    a.causal, a.dtype = (1 if causal else 0), _dtype_code(q)
    h.pieces = int(chunks) if chunks else 0
    with torch.cuda.device(dev):
        check(lib.sfa_fwd_host(h, _stream()))
    return out, lse


def fwd_host_plan(batch: int, seqlen_q: int, seqlen_k: int, heads_q: int, heads_kv: int, head_dim: int,
                  causal: bool = True, chunks: Optional[int] = None):
    """The piece plan fwd_host follows for these dimensions: a list of (batch entry, first kv-head, kv-heads, first
    row, end row, unit) in issue order (sfa_fwd_host_plan; pure host logic, needs no GPU)."""
    import ctypes as C
    h = _capi.FwdHostArgs()
    a = h.t
    a.batch, a.seqlen_q, a.seqlen_k, a.heads_q, a.heads_kv, a.head_dim = batch, seqlen_q, seqlen_k, heads_q, heads_kv, head_dim
    a.causal = 1 if causal else 0
    h.pieces = int(chunks) if chunks else 0
    n = lib.sfa_fwd_host_plan(h, None, 0)
    if n < 0:
        check(n)
    buf = (C.c_int32 * (6 * max(n, 1)))()
    n = lib.sfa_fwd_host_plan(h, buf, n)
    return [tuple(buf[6 * i:6 * i + 6]) for i in range(n)]


def fwd_host_last_pieces() -> int:
    """Diagnostics: number of pieces the most recent fwd_host call was cut into."""
    return int(lib.sfa_fwd_host_last_pieces())


def decode_host(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, seqlens: Optional[torch.Tensor] = None,
                softmax_scale: Optional[float] = None, out: Optional[torch.Tensor] = None,
                lse: Optional[torch.Tensor] = None, device=None, chunks: Optional[int] = None):
    """Split-KV decode on pinned HOST tensors: q (B,Hq,D), caches (B,M,Hkv,D), seqlens int32 (B) or None ->
    (out, lse) in pinned host memory.  The batch is cut into pieces whose cache upload overlaps the previous
    piece's kernel and download.  Asynchronous on the current stream of ``device``."""
    for t, n in ((q, "q"), (k_cache, "k_cache"), (v_cache, "v_cache")):
        _pinned(t, n)
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    B, Hq, D = q.shape
    if seqlens is not None:
        _pinned(seqlens, "seqlens")
        if seqlens.dtype != torch.int32 or seqlens.numel() != B or not seqlens.is_contiguous():
            raise ValueError("seqlens must be a contiguous int32 tensor of shape (batch,)")
    if out is None:
        out = torch.empty(q.shape, dtype=q.dtype, pin_memory=True)
    if lse is None:
        lse = torch.empty((B, Hq), dtype=torch.float32, pin_memory=True)
    _pinned(out, "out"); _pinned(lse, "lse")
    if not lse.is_contiguous() or lse.dtype != torch.float32:
        raise ValueError("lse must be a contiguous float32 (batch, heads_q) tensor")
    h = _capi.DecodeHostArgs()
    h.t = _decode_args(q, k_cache, v_cache, seqlens, out, lse, softmax_scale, 0)
    h.pieces = int(chunks) if chunks else 0
    with torch.cuda.device(dev):
        check(lib.sfa_decode_host(h, _stream()))
    return out, lse
