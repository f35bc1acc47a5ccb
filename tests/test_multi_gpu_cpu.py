"""CPU: the batch x kv-head sharding plan, and the N>1 path over gloo with world_size 2 (the oracle
stands in for the kernels; only the host logic is under test here)."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_partition_is_contiguous_balanced_and_complete():
    from starflashattention_b200.multi_gpu import partition
    for n in (0, 1, 7, 8, 64, 257):
        for w in (1, 2, 3, 4, 8):
            parts = [partition(n, w, r) for r in range(w)]
            assert sum(len(p) for p in parts) == n
            assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1
            flat = [u for p in parts for u in p]
            assert flat == list(range(n))
    with pytest.raises(ValueError):
        partition(4, 2, 2)


def test_shard_batch_heads_keeps_groups_whole():
    from starflashattention_b200.multi_gpu import shard_batch_heads
    for (B, Hkv, W) in [(8, 8, 8), (8, 8, 3), (1, 32, 8), (3, 5, 4), (2, 1, 4)]:
        seen = []
        for r in range(W):
            for seg in shard_batch_heads(B, Hkv, W, r):
                assert 0 <= seg.head_begin < seg.head_end <= Hkv
                seen += [(seg.batch, h) for h in range(seg.head_begin, seg.head_end)]
        assert seen == [(b, h) for b in range(B) for h in range(Hkv)]


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import naive
        from starflashattention_b200 import multi_gpu
        torch.manual_seed(0)
        B, S, Hq, Hkv, D = 3, 40, 6, 3, 16
        qq, kk, vv = torch.randn(B, S, Hq, D), torch.randn(B, S, Hkv, D), torch.randn(B, S, Hkv, D)
        g = Hq // Hkv
        results = []
        for seg in multi_gpu.shard_batch_heads(B, Hkv, world, rank):
            o, lse = naive.attention_fwd(qq[seg.batch:seg.batch + 1, :, seg.head_begin * g:seg.head_end * g],
                                         kk[seg.batch:seg.batch + 1, :, seg.head_begin:seg.head_end],
                                         vv[seg.batch:seg.batch + 1, :, seg.head_begin:seg.head_end], causal=True)
            results.append((seg, o, lse))
        local = multi_gpu.assemble(results, B, Hq, Hkv, (S, D), qq, head_axis=2)
        full = multi_gpu.gather_for_verification(local)
        ref, _ = naive.attention_fwd(qq, kk, vv, causal=True)
        q.put((rank, float((full - ref).abs().max()), len(results)))
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo_shards_and_gathers():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, err, nseg in out:
        assert err < 1e-5 and nseg >= 1


# --------------------------------------------------------------------------- KV-length sharding (row f2)
def test_kv_shard_range_covers_the_cache_in_multiples_of_64():
    from starflashattention_b200.multi_gpu import kv_shard_range, local_seqlens
    for total in (1, 63, 64, 1000, 32768, 65536 + 5):
        for w in (1, 2, 3, 4, 8):
            shards = [kv_shard_range(total, w, r) for r in range(w)]
            flat = [shards[0].start] + [s.stop for s in shards]
            assert shards[0].start == 0 and shards[-1].stop == total
            assert all(a.stop == b.start for a, b in zip(shards, shards[1:]))
            assert all(len(s) % 64 == 0 for s in shards if s.stop != total)
            sl = torch.tensor([0, 1, total // 2, total], dtype=torch.int32)
            loc = torch.stack([local_seqlens(sl, s) for s in shards])
            assert torch.equal(loc.sum(0).to(torch.int32), sl)        # every visible key is counted exactly once
            assert flat == sorted(flat)


def _kv_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import naive
        from starflashattention_b200 import multi_gpu
        torch.manual_seed(0)
        B, M, Hq, Hkv, D = 2, 333, 4, 2, 16
        qq, kk, vv = torch.randn(B, Hq, D), torch.randn(B, M, Hkv, D), torch.randn(B, M, Hkv, D)
        seqlens = torch.tensor([333, 70], dtype=torch.int32)          # the second sequence ends inside shard 0
        shard = multi_gpu.kv_shard_range(M, world, rank)
        loc = multi_gpu.local_seqlens(seqlens, shard)
        o_r, lse_r = naive.decode_attention(qq, kk[:, shard.start:shard.stop], vv[:, shard.start:shard.stop], loc)
        # the exchange step: every rank ends up with every rank's (O_r, LSE_r) rows
        o_all = [torch.empty_like(o_r) for _ in range(world)]
        l_all = [torch.empty_like(lse_r) for _ in range(world)]
        dist.all_gather(o_all, o_r)
        dist.all_gather(l_all, lse_r)
        o, lse = naive.combine_splits(torch.stack(o_all, dim=-2), torch.stack(l_all, dim=-1),
                                      torch.ones(B, Hq, world))
        ref_o, ref_lse = naive.decode_attention(qq, kk, vv, seqlens)
        q.put((rank, float((o - ref_o).abs().max()), float((lse - ref_lse).abs().max())))
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo_kv_length_sharded_decode_merges_to_the_full_result():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + os.getpid() % 2000
    procs = [ctx.Process(target=_kv_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, err_o, err_lse in out:
        assert err_o < 1e-5 and err_lse < 1e-5
