"""GPU, two or more devices (skipped on a one-GPU box): the multi-GPU driver with REAL processes, one per GPU.

* KVShardedDecode: the KV length of the sequences is cut over the ranks; `exchange="peer"` stores (O, LSE) rows into
  the peers' CUDA-IPC mapped workspaces over NVLink inside the merge kernel (st.release.sys / ld.acquire.sys across
  devices - the memory-model situation the one-GPU cooperative emulation of test_shard_gpu.py cannot show),
  `exchange="nccl"` is publish + all_gather + merge.  Both against the CPU oracle on the full cache.
* shard_batch_heads / sharded_fwd: the (batch x kv-head) units of one prefill problem cut over the ranks, gathered for
  verification, against the oracle.
The reference has no multi-GPU code; the contract is starflashattention_b200/multi_gpu.py."""
import os
import socket

import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, what, ret):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        import starflashattention_b200 as sfa
        from starflashattention_b200 import multi_gpu
        from oracle import naive
        g = torch.Generator().manual_seed(5)                       # the same inputs on every rank
        if what in ("peer", "nccl"):
            dtype = torch.bfloat16
            B, Hq, Hkv, D, M = 2, 32, 8, 128, 6000
            q = torch.randn(B, Hq, D, generator=g).to(dtype)
            kc = torch.randn(B, M, Hkv, D, generator=g).to(dtype)
            vc = torch.randn(B, M, Hkv, D, generator=g).to(dtype)
            seqlens = torch.tensor([6000, 2500], dtype=torch.int32)
            shard = multi_gpu.kv_shard_range(M, world, rank)
            dec = multi_gpu.KVShardedDecode(B, Hq, Hkv, D, len(shard), rank, world, dev, exchange=what)
            k_sh, v_sh = kc[:, shard.start:shard.stop].to(dev), vc[:, shard.start:shard.stop].to(dev)
            loc = multi_gpu.local_seqlens(seqlens, shard).to(dev)
            o_ref, lse_ref = naive.decode_attention(q, kc, vc, seqlens)
            worst = 0.0
            for step in range(5):                                   # both parities of the double buffer, reuse
                o, lse = dec.decode(q.to(dev), k_sh, v_sh, loc)
                torch.cuda.synchronize()
                err = (o.float().cpu() - o_ref).abs()
                # bf16 output: half an ulp of O (2^-9 relative) on top of north_star's 2e-3
                assert bool((err <= 2e-3 + 2.0 ** -8 * o_ref.abs()).all()), (what, rank, step, float(err.max()))
                assert float((lse.cpu() - lse_ref).abs().max()) <= 1e-3
                worst = max(worst, float(err.max()))
            dec.check()
            dec.close()
            ret[rank] = worst
        else:                                                       # units of one prefill problem cut over the ranks
            dtype = torch.float16
            B, S, Hq, Hkv, D = 2, 640, 8, 4, 128
            q = torch.randn(B, S, Hq, D, generator=g).to(dtype)
            k = torch.randn(B, S, Hkv, D, generator=g).to(dtype)
            v = torch.randn(B, S, Hkv, D, generator=g).to(dtype)
            res = multi_gpu.sharded_fwd(q.to(dev), k.to(dev), v.to(dev), world, rank, causal=True)
            full = multi_gpu.assemble(res, B, Hq, Hkv, (S, D), q.to(dev), head_axis=2)
            full = multi_gpu.gather_for_verification(full)
            torch.cuda.synchronize()
            assert sfa.fwd_watchdog() == 0
            o_ref, _ = naive.attention_fwd(q, k, v, causal=True)
            err = float((full.float().cpu() - o_ref).abs().max())
            assert err <= 2e-3, err
            ret[rank] = err
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("what", ["peer", "nccl", "fwd_units"])
def test_two_processes_two_gpus(what):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (run with gpurun --gpus 2)")
    import torch.multiprocessing as mp
    world = 2
    ret = mp.get_context("spawn").Manager().dict()
    mp.spawn(_worker, args=(world, _free_port(), what, ret), nprocs=world, join=True)
    assert len(ret) == world
