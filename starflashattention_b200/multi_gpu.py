"""Multi-GPU driver: shard the (batch x kv-head) space over the GPUs of one box.

Every (batch, kv-head group) of an attention call is independent, so the
path partitions with no exchange step: one process per GPU (torchrun), each rank runs the
single-GPU kernels on its own contiguous slice of units, and nothing crosses NVLink inside the
timed path.  ``gather_for_verification`` is the only collective and exists for tests.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional, Tuple

import torch


def partition(n_units: int, world_size: int, rank: int) -> range:
    """Contiguous, balanced slice of ``range(n_units)`` owned by ``rank`` (sizes differ by at most one)."""
    if world_size <= 0 or not (0 <= rank < world_size):
        raise ValueError("bad world_size / rank")
    base, extra = divmod(n_units, world_size)
    start = rank * base + min(rank, extra)
    return range(start, start + base + (1 if rank < extra else 0))


@dataclass(frozen=True)
class Segment:
    """kv-heads [head_begin, head_end) of one batch entry."""
    batch: int
    head_begin: int
    head_end: int


def shard_batch_heads(batch: int, heads_kv: int, world_size: int, rank: int) -> List[Segment]:
    """Units are (batch, kv-head) pairs, batch-major; a rank's slice is cut into per-batch head ranges.
    A kv-head always stays with its heads_q/heads_kv query heads."""
    units = partition(batch * heads_kv, world_size, rank)
    segs: List[Segment] = []
    u = units.start
    while u < units.stop:
        b, h0 = divmod(u, heads_kv)
        h1 = min(heads_kv, h0 + (units.stop - u))
        segs.append(Segment(b, h0, h1))
        u += h1 - h0
    return segs


def sharded_fwd(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, world_size: int, rank: int,
                causal: bool = False, softmax_scale: Optional[float] = None):
    """Runs this rank's share of ``fwd`` on full-size q/k/v views.  Returns [(segment, out, lse)]."""
    from .interface import fwd
    B, _, Hq, _ = q.shape
    Hkv = k.shape[2]
    g = Hq // Hkv
    results = []
    for seg in shard_batch_heads(B, Hkv, world_size, rank):
        qs = q[seg.batch:seg.batch + 1, :, seg.head_begin * g:seg.head_end * g]
        ks = k[seg.batch:seg.batch + 1, :, seg.head_begin:seg.head_end]
        vs = v[seg.batch:seg.batch + 1, :, seg.head_begin:seg.head_end]
        out, lse = fwd(qs, ks, vs, causal=causal, softmax_scale=softmax_scale)
        results.append((seg, out, lse))
    return results


def sharded_decode(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, seqlens: Optional[torch.Tensor],
                   world_size: int, rank: int, softmax_scale: Optional[float] = None, num_splits: int = 0):
    """Runs this rank's share of ``decode``.  Returns [(segment, out, lse)]."""
    from .interface import decode
    B, Hq, _ = q.shape
    Hkv = k_cache.shape[2]
    g = Hq // Hkv
    results = []
    for seg in shard_batch_heads(B, Hkv, world_size, rank):
        qs = q[seg.batch:seg.batch + 1, seg.head_begin * g:seg.head_end * g]
        ks = k_cache[seg.batch:seg.batch + 1, :, seg.head_begin:seg.head_end]
        vs = v_cache[seg.batch:seg.batch + 1, :, seg.head_begin:seg.head_end]
        sl = seqlens[seg.batch:seg.batch + 1] if seqlens is not None else None
        out, lse = decode(qs, ks, vs, sl, softmax_scale=softmax_scale, num_splits=num_splits)
        results.append((seg, out, lse))
    return results


def assemble(results, batch: int, heads_q: int, heads_kv: int, tail_shape: Tuple[int, ...], like: torch.Tensor,
             head_axis: int) -> torch.Tensor:
    """Places per-segment outputs into a zero-initialised full tensor (for verification)."""
    g = heads_q // heads_kv
    if head_axis == 2:      # prefill out: (B, S, Hq, D)
        full = torch.zeros((batch, tail_shape[0], heads_q, tail_shape[1]), dtype=like.dtype, device=like.device)
        for seg, out, _ in results:
            full[seg.batch, :, seg.head_begin * g:seg.head_end * g] = out[0]
    else:                   # decode out: (B, Hq, D)
        full = torch.zeros((batch, heads_q, tail_shape[0]), dtype=like.dtype, device=like.device)
        for seg, out, _ in results:
            full[seg.batch, seg.head_begin * g:seg.head_end * g] = out[0]
    return full


def gather_for_verification(local_full: torch.Tensor) -> torch.Tensor:
    """Sum of the ranks' disjoint, zero-padded outputs (all_reduce; outside any timed region)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        t = local_full.float()
        dist.all_reduce(t)
        return t.to(local_full.dtype)
    return local_full
