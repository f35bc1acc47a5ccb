"""Cross-GPU split-KV decode of ONE long sequence, one process per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
        tools/shard_decode_bench.py [kv_len] [iters]

Every rank holds 1/N of the KV rows.  Timed per step on the device (CUDA events, max over ranks): the fused peer
exchange (P2P stores + flags inside the merge kernel) and the NCCL form (publish kernel, all_gather, merge
kernel).  Rank 0 also checks the result against the single-GPU kernel over the whole cache when it fits.
Prints one JSON line."""
import json
import os
import statistics
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import starflashattention_b200 as sfa  # noqa: E402
from starflashattention_b200 import multi_gpu  # noqa: E402


def main():
    M = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 30
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B, Hq, Hkv, D, dt = 1, 32, 8, 128, torch.bfloat16
    shard = multi_gpu.kv_shard_range(M, world, rank)
    g = torch.Generator(device=dev).manual_seed(1234)            # same stream on every rank: slice the same cache
    q = torch.randn((B, Hq, D), generator=g, device=dev).to(dt)
    # generate the shard deterministically per rank (a full 256K-row cache per rank would waste memory)
    gk = torch.Generator(device=dev).manual_seed(100 + rank)
    ks = torch.randn((B, len(shard), Hkv, D), generator=gk, device=dev).to(dt)
    vs = torch.randn((B, len(shard), Hkv, D), generator=gk, device=dev).to(dt)
    seqlens = torch.tensor([M], dtype=torch.int32)
    loc = multi_gpu.local_seqlens(seqlens, shard).to(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    res = {}
    outs = {}
    for mode in (("peer", "nccl") if world > 1 else ("peer",)):
        dec = multi_gpu.KVShardedDecode(B, Hq, Hkv, D, len(shard), rank, world, dev, exchange=mode)
        for _ in range(3):
            o, lse = dec.decode(q, ks, vs, loc)
        torch.cuda.synchronize()
        ts = []
        out_b, lse_b = torch.empty_like(q), torch.empty((B, Hq), dtype=torch.float32, device=dev)
        inner = 8                      # steps per timed region: the host work of a call hides behind the previous step
        for _ in range(iters):
            flush.zero_()
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(inner):
                o, lse = dec.decode(q, ks, vs, loc, out=out_b, lse=lse_b)
            e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1) / inner], device=dev)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ts.append(float(t) * 1e3)
        wd = dec.watchdog()
        res[mode] = {"us_median": statistics.median(ts), "us_min": min(ts), "watchdog": wd,
                     "splits_per_rank": dec.num_splits}
        outs[mode] = (o.float().clone(), lse.clone())
        dec.close()
    # correctness: gather the shards on rank 0 and run the single-GPU kernel over the whole cache
    check = None
    if world > 1:
        k_all = [torch.empty_like(ks) for _ in range(world)] if rank == 0 else None
        v_all = [torch.empty_like(vs) for _ in range(world)] if rank == 0 else None
        same = len(shard) * world == M
        if same:
            dist.gather(ks, k_all, dst=0)
            dist.gather(vs, v_all, dst=0)
            if rank == 0:
                o_ref, lse_ref = sfa.decode(q, torch.cat(k_all, 1), torch.cat(v_all, 1))
                torch.cuda.synchronize()
                check = {m: {"o_max_abs_vs_single_gpu": float((outs[m][0] - o_ref.float()).abs().max()),
                             "lse_max_abs": float((outs[m][1] - lse_ref).abs().max())} for m in outs}
    else:
        o_ref, lse_ref = sfa.decode(q, ks, vs)
        torch.cuda.synchronize()
        check = {"peer": {"o_max_abs_vs_single_gpu": float((outs["peer"][0] - o_ref.float()).abs().max()),
                          "lse_max_abs": float((outs["peer"][1] - lse_ref).abs().max())}}
    if rank == 0:
        kv_bytes = 2.0 * B * M * Hkv * D * 2
        line = {"metric": "kv_sharded_decode_us_per_step", "n_gpus": world, "kv_len": M, "heads": [Hq, Hkv], "head_dim": D,
                "dtype": "bf16", "kv_bytes": kv_bytes, "exchange_bytes_per_rank": B * Hq * (D + 1) * 4,
                "timing": "CUDA events around 8 back-to-back steps (KV larger than L2 when sharded 1/N >= 126 MB), max over ranks, median of the regions", "check": check}
        for m, r in res.items():
            line[m] = dict(r, aggregate_gbs=kv_bytes / r["us_median"] / 1e3)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
