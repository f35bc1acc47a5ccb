"""Multi-GPU driver: shard the (batch x kv-head) space over the GPUs of one box.

Every (batch, kv-head group) of an attention call is independent, so the
path partitions with no exchange step: one process per GPU (torchrun), each rank runs the
single-GPU kernels on its own contiguous slice of units, and nothing crosses NVLink inside the
timed path.  ``gather_for_verification`` is the only collective and exists for tests.

The one place with a real exchange step is a single long sequence whose KV length is cut over the GPUs
(``KVShardedDecode``): each rank attends over its shard and the ranks swap
(O, LSE) rows - 16.5 KB per rank at C4's head shape - through peer memory inside the merge kernel.
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


# --------------------------------------------------------------------------- KV-length sharding (row f2)
def kv_shard_range(total_len: int, world_size: int, rank: int, multiple: int = 64) -> range:
    """Keys [begin, end) of a cache of ``total_len`` rows owned by ``rank``: equal shards rounded up to a
    multiple of ``multiple`` rows (the last ranks may own fewer rows, or none)."""
    if world_size <= 0 or not (0 <= rank < world_size):
        raise ValueError("bad world_size / rank")
    shard = -(-total_len // world_size)
    shard = -(-shard // multiple) * multiple
    begin = min(total_len, rank * shard)
    return range(begin, min(total_len, begin + shard))


def local_seqlens(seqlens: torch.Tensor, shard: range) -> torch.Tensor:
    """Visible keys of every sequence inside ``shard`` (global lengths clipped to the shard)."""
    return (seqlens.to(torch.int64) - shard.start).clamp_(0, len(shard)).to(torch.int32)


class KVShardedDecode:
    """Split-KV decode of sequences whose KV length is sharded over ``world`` GPUs, one process per GPU.

    ``exchange="nccl"`` (default): publish kernel, NCCL ``all_gather_into_tensor``, merge kernel.
    ``exchange="peer"`` (opt-in): the merge kernel stores this rank's (O, LSE) rows into every peer's workspace over
    NVLink and merges the world's rows in the same launch (CUDA-IPC mapped buffers; ``torch.distributed`` is
    used once, to hand the 64-byte IPC handles around).  It is the faster form (one launch, no NCCL call), but its
    blocks spin on the peers' flags: every rank must make the same calls in the same order, ``batch * heads_q``
    must fit one resident wave (checked by the library), and a peer that never arrives turns the affected rows
    into NaN after 2 s and latches :meth:`watchdog` - call :meth:`check` at a point where a device synchronise
    is acceptable; after a failure the instance refuses further steps (the ranks' step counters no longer agree).
    ``peer_ws`` lets a test emulate several ranks inside one process by passing the workspaces directly."""

    def __init__(self, batch: int, heads_q: int, heads_kv: int, head_dim: int, shard_len: int, rank: int, world: int,
                 device, exchange: str = "nccl", group=None, peer_ws: Optional[List[int]] = None):
        from . import _capi
        if exchange not in ("peer", "nccl"):
            raise ValueError("exchange must be 'peer' or 'nccl'")
        self._capi, self.lib = _capi, _capi.lib
        self.batch, self.heads_q, self.heads_kv, self.head_dim = batch, heads_q, heads_kv, head_dim
        self.rank, self.world, self.exchange, self.group = rank, world, exchange, group
        self.device = torch.device(device)
        self.step = 0
        self._failed = 0
        self._owned = None
        self._opened: List[int] = []
        rows = batch * heads_q
        with torch.cuda.device(self.device):
            a = _capi.DecodeArgs()
            a.batch, a.seqlen_k, a.heads_kv, a.num_splits = batch, max(1, shard_len), heads_kv, 0
            self.num_splits = max(1, int(self.lib.sfa_decode_num_splits(a)))
            part = rows * self.num_splits
            self.workspace = torch.empty(part * head_dim * 4 + part * 8 + 512, dtype=torch.uint8, device=self.device)
            if exchange == "nccl":
                self.payload = torch.empty((rows * (head_dim + 1),), dtype=torch.float32, device=self.device)
                self.gathered = torch.empty((world, rows * (head_dim + 1)), dtype=torch.float32, device=self.device)
                self.peers = None
            elif peer_ws is not None:
                self.peers = list(peer_ws)
            else:
                import ctypes as C
                import torch.distributed as dist
                nbytes = int(self.lib.sfa_shard_workspace_bytes(world, batch, heads_q, head_dim))
                ptr, handle = C.c_void_p(), C.create_string_buffer(64)
                _capi.check(self.lib.sfa_peer_alloc(nbytes, C.byref(ptr), handle))
                self._owned = ptr.value
                handles: List[Optional[bytes]] = [handle.raw] * world
                if world > 1:
                    dist.all_gather_object(handles, handle.raw, group=group)
                self.peers = []
                for r in range(world):
                    if r == rank:
                        self.peers.append(self._owned)
                    else:
                        p = C.c_void_p()
                        _capi.check(self.lib.sfa_peer_open(handles[r], C.byref(p)))
                        self._opened.append(p.value)
                        self.peers.append(p.value)
                if world > 1:
                    dist.barrier(group=group)

    @staticmethod
    def workspace_bytes(world: int, batch: int, heads_q: int, head_dim: int) -> int:
        from . import _capi
        return int(_capi.lib.sfa_shard_workspace_bytes(world, batch, heads_q, head_dim))

    def decode(self, q: torch.Tensor, k_shard: torch.Tensor, v_shard: torch.Tensor, seqlens_local: torch.Tensor,
               out: Optional[torch.Tensor] = None, lse: Optional[torch.Tensor] = None,
               softmax_scale: Optional[float] = None):
        """q (B, Hq, D); k_shard / v_shard (B, shard_len, Hkv, D): this rank's rows of the cache; seqlens_local
        int32 (B): visible keys inside the shard.  Returns the full (out, lse) on every rank, asynchronously on
        the current stream."""
        from .interface import _decode_args, _stream
        capi = self._capi
        if self._failed:
            raise RuntimeError(f"KVShardedDecode: rank {self._failed - 1} did not arrive in an earlier step; the ranks' "
                               "step counters no longer agree - build a new instance on every rank")
        if out is None:
            out = torch.empty_like(q)
        if lse is None:
            lse = torch.empty(q.shape[:2], dtype=torch.float32, device=q.device)
        a = _decode_args(q, k_shard, v_shard, seqlens_local, out, lse, softmax_scale, self.num_splits)
        a.workspace, a.workspace_bytes = self.workspace.data_ptr(), self.workspace.numel()
        self.step += 1
        c = capi.ShardComm()
        c.rank, c.world, c.step = self.rank, self.world, self.step
        with torch.cuda.device(self.device):
            if self.exchange == "peer":
                c.mode = 0
                for r in range(self.world):
                    c.peer_ws[r] = self.peers[r]
                capi.check(self.lib.sfa_decode_sharded(a, c, _stream()))
            else:
                import torch.distributed as dist
                c.mode, c.payload = 1, self.payload.data_ptr()
                capi.check(self.lib.sfa_decode_sharded(a, c, _stream()))
                dist.all_gather_into_tensor(self.gathered, self.payload, group=self.group)
                capi.check(self.lib.sfa_shard_merge(self.gathered.data_ptr(), self.world, out.data_ptr(), lse.data_ptr(),
                                                    q.shape[0], q.shape[1], q.shape[2], out.stride(0), out.stride(1),
                                                    a.dtype, _stream()))
        return out, lse

    def watchdog(self, clear: bool = True) -> int:
        """Synchronises the device; non-zero (1 + the missing rank) after a peer's rows did not arrive within 2 s."""
        with torch.cuda.device(self.device):
            rc = int(self.lib.sfa_shard_watchdog(1 if clear else 0))
        if rc > 0:
            self._failed = rc
        return rc

    def check(self) -> None:
        """Raises if an exchange timed out since the last check (the affected output rows are NaN)."""
        rc = self.watchdog()
        if rc > 0:
            raise RuntimeError(f"KVShardedDecode: rank {rc - 1}'s rows did not arrive within 2 s; the outputs of that "
                               "step are NaN")
        if rc < 0:
            self._capi.check(rc)

    def close(self):
        """Unmaps the peers' buffers and frees the own one (collective: every rank calls it)."""
        if self._owned is None:
            return
        torch.cuda.synchronize(self.device)
        for p in self._opened:
            self.lib.sfa_peer_close(p)
        self._opened = []
        if self.world > 1:
            import torch.distributed as dist
            if dist.is_initialized():
                dist.barrier(group=self.group)
        self.lib.sfa_peer_free(self._owned)
        self._owned = None
