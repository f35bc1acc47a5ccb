"""The box's own ceiling for the end-to-end leg: plain pinned cudaMemcpyAsync, host -> device and device -> host at
the same time on two streams, on every rank at once - no kernel, no pipeline.  Run it with 1, 2, 4 and 8 processes:

    python tools/pcie_ceiling.py                                  (one GPU)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/pcie_ceiling.py

Per rank it moves the bytes of its share of C3 (2.16 GB / N) and of C5 (2.16 GB / N) and, as a reference point, 1 GiB
each way; prints one JSON line with the per-rank time (max over ranks) and the aggregate GB/s of the box."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


def main():
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    cases = {"c3_share": (3 * (1 << 29) // world, ((1 << 29) + (8 << 20)) // world),      # q, k, v up; o + lse down
             "c5_share": (3 * (1 << 29) // world, ((1 << 29) + (8 << 20)) // world),
             "1GiB_each_way": (1 << 30, 1 << 30), "1GiB_up_only": (1 << 30, 0), "1GiB_down_only": (0, 1 << 30)}
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
    res = {}
    for name, (up, down) in cases.items():
        h_up = torch.empty(max(up, 1), dtype=torch.uint8, pin_memory=True)
        d_up = torch.empty(max(up, 1), dtype=torch.uint8, device=dev)
        h_dn = torch.empty(max(down, 1), dtype=torch.uint8, pin_memory=True)
        d_dn = torch.empty(max(down, 1), dtype=torch.uint8, device=dev)
        best = float("inf")
        for _ in range(4):
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            cur = torch.cuda.current_stream()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(cur)
            s_in.wait_event(e0); s_out.wait_event(e0)
            if up:
                with torch.cuda.stream(s_in):
                    d_up.copy_(h_up, non_blocking=True)
            if down:
                with torch.cuda.stream(s_out):
                    h_dn.copy_(d_dn, non_blocking=True)
            cur.wait_stream(s_in); cur.wait_stream(s_out)
            e1.record(cur)
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        t = torch.tensor([best], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        res[name] = {"ms": ms, "up_bytes_per_rank": up, "down_bytes_per_rank": down,
                     "aggregate_up_gbs": up * world / ms / 1e6, "aggregate_down_gbs": down * world / ms / 1e6}
        del h_up, d_up, h_dn, d_dn
    if rank == 0:
        print(json.dumps({"n_gpus": world, "host_cpus": os.cpu_count(), "copies": res}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
