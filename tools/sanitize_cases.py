"""Tiny invocations of every kernel family for compute-sanitizer (memcheck / racecheck / synccheck / initcheck):

    compute-sanitizer --tool memcheck python tools/sanitize_cases.py fwd decode step shard

Each case is checked against the CPU oracle as well, so a sanitizer-clean run is also a correct run.
"""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import starflashattention_b200 as sfa  # noqa: E402
from oracle import naive  # noqa: E402


def rnd(shape, dtype, seed):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(shape, generator=g).to(dtype)


def case_fwd():
    for (B, Sq, Sk, H, Hkv, D, causal, dt) in [(1, 256, 256, 2, 2, 128, True, torch.bfloat16), (1, 200, 333, 2, 1, 128, False, torch.float16),
                                               (2, 640, 640, 2, 2, 128, True, torch.bfloat16), (1, 300, 300, 2, 2, 64, True, torch.float16),
                                               (1, 100, 419, 2, 2, 128, True, torch.float16), (40, 700, 700, 4, 2, 128, True, torch.bfloat16)]:
        q, k, v = rnd((B, Sq, H, D), dt, 1), rnd((B, Sk, Hkv, D), dt, 2), rnd((B, Sk, Hkv, D), dt, 3)
        o, lse = sfa.fwd(q.cuda(), k.cuda(), v.cuda(), causal=causal)
        torch.cuda.synchronize()
        o_ref, lse_ref = naive.attention_fwd(q, k, v, causal=causal)
        err = float((o.float().cpu() - o_ref).abs().max())
        assert err < 2e-2 and sfa.fwd_watchdog() == 0, err
        print(f"[sanitize] fwd {B}x{Sq}x{Sk}x{H}x{D} causal={causal}: max err {err:.2e}")


def case_decode():
    dt = torch.bfloat16
    q, kc, vc = rnd((2, 8, 128), dt, 4), rnd((2, 1000, 2, 128), dt, 5), rnd((2, 1000, 2, 128), dt, 6)
    sl = torch.tensor([1000, 333], dtype=torch.int32)
    o_ref, _ = naive.decode_attention(q, kc, vc, sl)
    for splits in (8, 12, 1):      # cluster merge, two-launch form, direct write
        o, _ = sfa.decode(q.cuda(), kc.cuda(), vc.cuda(), sl.cuda(), num_splits=splits)
        torch.cuda.synchronize()
        err = float((o.float().cpu() - o_ref).abs().max())
        assert err < 2e-2, err
        print(f"[sanitize] decode {splits} splits: max err {err:.2e}")


def case_step():
    dt = torch.float16
    B, L, M, H, D = 2, 2, 256, 4, 128
    qkv, kt, vt = rnd((B, 3, H, D), dt, 7), rnd((B, L, M, H, D), dt, 8), rnd((B, L, M, H, D), dt, 9)
    seq = torch.tensor([100, 255], dtype=torch.int32)
    cos, sin = sfa.compute_rotary_table(M, D, dtype=dt)
    o_ref, *_ = naive.decode_step_reference(qkv, kt, vt, seq, 1, table_dtype=dt)
    o = torch.empty(B, H, D, dtype=dt, device="cuda")
    sfa.mha_fwd_cuda(qkv.cuda(), None, None, None, kt.cuda(), vt.cuda(), seq.cuda(), o, B, M, H, D, D, M, L, 1,
                     rotary_cos_table=cos, rotary_sin_table=sin, num_splits=4)
    torch.cuda.synchronize()
    err = float((o.float().cpu() - o_ref).abs().max())
    assert err < 1e-2, err
    print(f"[sanitize] decoder step (fused RoPE + append, 4 cluster splits, tables): max err {err:.2e}")
    # cache-full with tables: position == memory_max_len must not read a table row past the end
    seq_full = torch.tensor([M, 10], dtype=torch.int32)
    sfa.mha_fwd_cuda(qkv.cuda(), None, None, None, kt.cuda(), vt.cuda(), seq_full.cuda(), o, B, M, H, D, D, M, L, 1,
                     rotary_cos_table=cos, rotary_sin_table=sin, num_splits=2)
    torch.cuda.synchronize()
    print("[sanitize] decoder step, cache full with tables: ran")
    # paged step, GQA
    page, pages, Hq, Hkv = 16, 40, 8, 2
    kp, vp = rnd((pages, page, Hkv, D), dt, 10), rnd((pages, page, Hkv, D), dt, 11)
    table = torch.randperm(pages)[:32].reshape(2, 16).to(torch.int32)
    qkv2 = rnd((2, Hq + 2 * Hkv, D), dt, 12)
    seq2 = torch.tensor([37, 200], dtype=torch.int32)
    o2 = sfa.decoder_step_paged(qkv2.cuda(), kp.cuda(), vp.cuda(), table.cuda(), seq2.cuda(), Hq, num_splits=3)
    torch.cuda.synchronize()
    assert torch.isfinite(o2).all()
    print("[sanitize] paged decoder step: ran")


def case_shard():
    import ctypes as C
    from starflashattention_b200 import _capi, multi_gpu
    from starflashattention_b200.interface import _decode_args, _stream
    dt, world, B, Hq, Hkv, D, M = torch.bfloat16, 4, 1, 8, 2, 128, 2000
    q, kc, vc = rnd((B, Hq, D), dt, 13), rnd((B, M, Hkv, D), dt, 14), rnd((B, M, Hkv, D), dt, 15)
    sl = torch.tensor([2000], dtype=torch.int32)
    o_ref, _ = naive.decode_attention(q, kc, vc, sl)
    dev = torch.device("cuda:0")
    nbytes = multi_gpu.KVShardedDecode.workspace_bytes(world, B, Hq, D)
    ws = [torch.zeros(nbytes, dtype=torch.uint8, device=dev) for _ in range(world)]
    ptrs = (C.c_void_p * world)(*[w.data_ptr() for w in ws])
    qd, kd, vd = q.to(dev), kc.to(dev), vc.to(dev)
    shards = [multi_gpu.kv_shard_range(M, world, r) for r in range(world)]
    locs = [multi_gpu.local_seqlens(sl, sh).to(dev) for sh in shards]
    part = [torch.empty(B * Hq * 3 * (D * 4 + 8) + 512, dtype=torch.uint8, device=dev) for _ in range(world)]
    for step in (1, 2):
        outs = [(torch.empty_like(qd), torch.empty((B, Hq), dtype=torch.float32, device=dev)) for _ in range(world)]
        args = (_capi.DecodeArgs * world)()
        for r, sh in enumerate(shards):
            a = _decode_args(qd, kd[:, sh.start:sh.stop], vd[:, sh.start:sh.stop], locs[r], outs[r][0], outs[r][1], None, 3)
            a.workspace, a.workspace_bytes = part[r].data_ptr(), part[r].numel()
            args[r] = a
        _capi.check(_capi.lib.sfa_decode_sharded_emulated(args, world, ptrs, step, _stream()))
        torch.cuda.synchronize()
        err = max(float((o.float().cpu() - o_ref).abs().max()) for o, _ in outs)
        assert err < 2e-2, err
    print(f"[sanitize] sharded decode, 4 emulated ranks, cooperative exchange: max err {err:.2e}")


if __name__ == "__main__":
    for name in sys.argv[1:] or ["fwd", "decode", "step", "shard"]:
        globals()["case_" + name]()
    print("[sanitize] done")
