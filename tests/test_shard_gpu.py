"""GPU parity of the cross-GPU split-KV decode on ONE device.  The ranks of a box
are emulated through sfa_decode_sharded_emulated: each rank's split-KV pass runs on its shard, then the exchange
blocks of ALL ranks go out as one cooperative launch whose workspaces stand in for the peers' CUDA-IPC mappings,
so the stores, flags, waits and merge run as they do over NVLink.  (Ranks as separate launches that spin on each
other on one GPU are not guaranteed to be co-resident - B200_PROFILING.md - so that form is not used.)  The real
multi-process run is tools/shard_decode_bench.py under torchrun (profiles/r1c_shard*_n*.json)."""
import pytest
import torch

from oracle import naive

pytestmark = pytest.mark.gpu


def _rand(shape, dtype, seed):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(shape, generator=g).to(dtype)


def _bound(mag, dtype):
    """north_star's 2e-3; bf16 adds the rounding of P (the tensor-core operand) and of O itself, each half an ulp =
    2^-9 relative, against the oracle's P @ |V| (>= |O|) on the same inputs - the bound of tests/test_fwd_gpu.py.  The
    cross-rank merge is fp32 and adds nothing at this scale."""
    return torch.full_like(mag, 2e-3) if dtype == torch.float16 else 2e-3 + 2.0 ** -8 * mag


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("world,B,Hq,Hkv,D,M,lens", [
    (2, 1, 32, 8, 128, 4096, [4096]),
    (4, 2, 8, 2, 128, 3000, [3000, 700]),        # the second sequence leaves the last shards empty
    (8, 1, 8, 8, 64, 8192, [8191]),
    (3, 2, 4, 4, 128, 1000, [1, 999]),
])
def test_peer_exchange_matches_oracle(dtype, world, B, Hq, Hkv, D, M, lens):
    import starflashattention_b200 as sfa
    from starflashattention_b200 import multi_gpu
    q = _rand((B, Hq, D), dtype, 1)
    kc = _rand((B, M, Hkv, D), dtype, 2)
    vc = _rand((B, M, Hkv, D), dtype, 3)
    seqlens = torch.tensor(lens, dtype=torch.int32)
    o_ref, lse_ref = naive.decode_attention(q, kc, vc, seqlens)
    mag, _ = naive.decode_attention(q, kc, vc.abs(), seqlens)
    import ctypes as C
    from starflashattention_b200 import _capi
    from starflashattention_b200.interface import _decode_args, _stream
    dev = torch.device("cuda:0")
    nbytes = multi_gpu.KVShardedDecode.workspace_bytes(world, B, Hq, D)
    ws = [torch.zeros(nbytes, dtype=torch.uint8, device=dev) for _ in range(world)]
    ws_ptrs = (C.c_void_p * world)(*[w.data_ptr() for w in ws])
    qd, kd, vd = q.to(dev), kc.to(dev), vc.to(dev)
    shards = [multi_gpu.kv_shard_range(M, world, r) for r in range(world)]
    locs = [multi_gpu.local_seqlens(seqlens, sh).to(dev) for sh in shards]
    splits = 5
    part = [torch.empty(B * Hq * splits * (D * 4 + 8) + 512, dtype=torch.uint8, device=dev) for _ in range(world)]
    for step in range(1, 4):                                # three steps: both parities of the double buffer, reuse
        outs = [(torch.empty_like(qd), torch.empty((B, Hq), dtype=torch.float32, device=dev)) for _ in range(world)]
        args = (_capi.DecodeArgs * world)()
        for r, sh in enumerate(shards):
            a = _decode_args(qd, kd[:, sh.start:sh.stop], vd[:, sh.start:sh.stop], locs[r], outs[r][0], outs[r][1],
                             None, splits)
            a.workspace, a.workspace_bytes = part[r].data_ptr(), part[r].numel()
            args[r] = a
        _capi.check(_capi.lib.sfa_decode_sharded_emulated(args, world, ws_ptrs, step, _stream()))
        torch.cuda.synchronize()
        assert _capi.lib.sfa_shard_watchdog(1) == 0, "a rank waited more than 2 s for a peer's rows"
        for o, lse in outs:                                 # every rank holds the full result
            err = (o.float().cpu() - o_ref).abs()
            assert bool((err <= _bound(mag, dtype)).all()), float(err.max())
            assert float((lse.cpu() - lse_ref).abs().max()) <= 1e-3
        assert all(torch.equal(outs[0][0], o) for o, _ in outs)


def test_publish_and_merge_form_matches_peer_form():
    """The NCCL baseline's two kernels (publish, merge of gathered payloads) with the all-gather done by hand."""
    import starflashattention_b200 as sfa
    from starflashattention_b200 import _capi, multi_gpu
    from starflashattention_b200.interface import _decode_args, _stream
    dtype, world, B, Hq, Hkv, D, M = torch.bfloat16, 4, 1, 16, 4, 128, 2048
    q, kc, vc = _rand((B, Hq, D), dtype, 4).cuda(), _rand((B, M, Hkv, D), dtype, 5).cuda(), _rand((B, M, Hkv, D), dtype, 6).cuda()
    seqlens = torch.tensor([2000], dtype=torch.int32)
    o_ref, lse_ref = naive.decode_attention(q.cpu(), kc.cpu(), vc.cpu(), seqlens)
    mag, _ = naive.decode_attention(q.cpu(), kc.cpu(), vc.cpu().abs(), seqlens)
    rows = B * Hq
    gathered = torch.empty((world, rows * (D + 1)), dtype=torch.float32, device="cuda")
    for r in range(world):
        shard = multi_gpu.kv_shard_range(M, world, r)
        out, lse = torch.empty_like(q), torch.empty((B, Hq), dtype=torch.float32, device="cuda")
        loc = multi_gpu.local_seqlens(seqlens, shard).cuda()
        a = _decode_args(q, kc[:, shard.start:shard.stop], vc[:, shard.start:shard.stop], loc, out, lse, None, 3)
        c = _capi.ShardComm()
        c.rank, c.world, c.step, c.mode, c.payload = r, world, 1, 1, gathered[r].data_ptr()
        _capi.check(_capi.lib.sfa_decode_sharded(a, c, _stream()))
    out, lse = torch.empty_like(q), torch.empty((B, Hq), dtype=torch.float32, device="cuda")
    _capi.check(_capi.lib.sfa_shard_merge(gathered.data_ptr(), world, out.data_ptr(), lse.data_ptr(), B, Hq, D,
                                          out.stride(0), out.stride(1), _capi.SFA_BF16, _stream()))
    torch.cuda.synchronize()
    err = (out.float().cpu() - o_ref).abs()
    assert bool((err <= _bound(mag, dtype)).all()), float(err.max())
    assert float((lse.cpu() - lse_ref).abs().max()) <= 1e-3


def test_sharded_decode_rejects_bad_arguments():
    from starflashattention_b200 import _capi
    a, c = _capi.DecodeArgs(), _capi.ShardComm()
    c.rank, c.world = 3, 2
    q = torch.zeros(1, 8, 128, dtype=torch.float16, device="cuda")
    kv = torch.zeros(1, 64, 8, 128, dtype=torch.float16, device="cuda")
    a.q, a.k_cache, a.v_cache, a.o = q.data_ptr(), kv.data_ptr(), kv.data_ptr(), q.data_ptr()
    a.batch, a.seqlen_k, a.heads_q, a.heads_kv, a.head_dim = 1, 64, 8, 8, 128
    assert _capi.lib.sfa_decode_sharded(a, c, None) == _capi.SFA_ERR_INVALID
    c.rank, c.world, c.mode = 0, 2, 0                        # peer mode without peer pointers
    assert _capi.lib.sfa_decode_sharded(a, c, None) == _capi.SFA_ERR_INVALID
