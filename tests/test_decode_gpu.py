"""GPU parity: split-KV decode and the drop-in decoder step, through the C ABI, against the CPU oracle
and the fixtures generated from the reference's own Python example.

Tolerances (north_star): |O - O_oracle| <= 2e-3 for fp16; for bf16 the same absolute bound plus the
output rounding of bf16 itself (half an ulp = 2^-9 relative), i.e. atol 2e-3 + rtol 4e-3.  LSE within 1e-3.
"""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import naive

pytestmark = pytest.mark.gpu

DT = {"float16": torch.float16, "bfloat16": torch.bfloat16, "float32": torch.float32}


def _tol(dtype):
    return (2e-3, 0.0) if dtype == torch.float16 else (2e-3, 4e-3)


def _assert_close(got, ref, dtype, what):
    atol, rtol = _tol(dtype)
    err = (got.float().cpu() - ref.float()).abs()
    bound = atol + rtol * ref.float().abs()
    assert bool((err <= bound).all()), f"{what}: max err {float(err.max()):.3e} (bound {float(bound.max()):.3e})"


def _rand(shape, dtype, seed, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(shape, generator=g) * scale).to(dtype)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("B,Hq,Hkv,D,M,lens,splits", [
    (2, 4, 4, 128, 300, [300, 129], 0),        # MHA, ragged, launcher's splits
    (1, 8, 2, 128, 1024, [1000], 1),           # GQA 4, single split (direct write)
    (3, 8, 1, 128, 700, [1, 64, 700], 5),      # GQA 8, one-key sequence, forced odd split count
    (2, 4, 2, 64, 513, [513, 17], 3),          # head_dim 64
    (2, 2, 2, 128, 256, [0, 256], 4),          # an empty sequence: O = 0, LSE = -inf
    (1, 32, 8, 128, 4096, [4096], 0),          # the C4 head layout at a CPU-checkable length
    (2, 32, 2, 128, 777, [777, 300], 0),       # GQA 16: two chunks of 8 query heads per kv head
    (1, 16, 1, 64, 1500, [1500], 6),           # MQA 16 at head_dim 64, cluster-sized split count
    (2, 24, 2, 128, 520, [100, 520], 3),       # GQA 12: two chunks of 6
    (1, 32, 1, 128, 2048, [2000], 0),          # MQA 32: four chunks of 8
    (1, 11, 1, 128, 300, [300], 2),            # a prime group above 8: eleven chunks of one head
])
def test_decode_matches_oracle(dtype, B, Hq, Hkv, D, M, lens, splits):
    import starflashattention_b200 as sfa
    q = _rand((B, Hq, D), dtype, 1)
    kc = _rand((B, M, Hkv, D), dtype, 2)
    vc = _rand((B, M, Hkv, D), dtype, 3)
    seqlens = torch.tensor(lens, dtype=torch.int32)
    o_ref, lse_ref = naive.decode_attention(q, kc, vc, seqlens)
    o, lse = sfa.decode(q.cuda(), kc.cuda(), vc.cuda(), seqlens.cuda(), num_splits=splits)
    torch.cuda.synchronize()
    _assert_close(o, o_ref, dtype, "O")
    lse = lse.cpu()
    fin = torch.isfinite(lse_ref)
    assert torch.equal(torch.isfinite(lse), fin)
    assert float((lse[fin] - lse_ref[fin]).abs().max() if fin.any() else 0.0) <= 1e-3


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_decode_untransposed_variant_matches_oracle(dtype, tuning):
    """decode_swap = 0: the untransposed MMA form of the split-KV kernel (query heads on the M rows), a rejected variant
    that only the -DSFA_PROBES library carries (SFA_PROBES_LIB=1 python -m pytest ... -k untransposed)."""
    import starflashattention_b200 as sfa
    if os.environ.get("SFA_PROBES_LIB") != "1":
        pytest.skip("the untransposed variant exists in the probe build only")
    tuning("decode_swap", 0, 1)
    B, Hq, Hkv, D, M = 2, 8, 2, 128, 1500
    q, kc, vc = _rand((B, Hq, D), dtype, 81), _rand((B, M, Hkv, D), dtype, 82), _rand((B, M, Hkv, D), dtype, 83)
    seqlens = torch.tensor([1500, 333], dtype=torch.int32)
    o_ref, lse_ref = naive.decode_attention(q, kc, vc, seqlens)
    o, lse = sfa.decode(q.cuda(), kc.cuda(), vc.cuda(), seqlens.cuda(), num_splits=3)
    torch.cuda.synchronize()
    _assert_close(o, o_ref, dtype, "O")
    assert float((lse.cpu() - lse_ref).abs().max()) <= 1e-3


def test_decode_strided_cache_and_no_seqlens():
    """Caches that are views into a (B, L, M, H, D) table (the drop-in layout), seqlens = None."""
    import starflashattention_b200 as sfa
    dtype = torch.bfloat16
    B, L, M, H, D = 2, 3, 200, 4, 128
    table_k = _rand((B, L, M, H, D), dtype, 5)
    table_v = _rand((B, L, M, H, D), dtype, 6)
    q = _rand((B, H, D), dtype, 7)
    o_ref, lse_ref = naive.decode_attention(q, table_k[:, 1], table_v[:, 1], torch.tensor([M, M]))
    tk, tv = table_k.cuda(), table_v.cuda()
    o, lse = sfa.decode(q.cuda(), tk[:, 1], tv[:, 1], None)
    _assert_close(o, o_ref, dtype, "O")
    assert float((lse.cpu() - lse_ref).abs().max()) <= 1e-3


def test_decode_split_counts_agree():
    """Size-independent property: the result does not depend on the number of splits."""
    import starflashattention_b200 as sfa
    dtype = torch.float16
    q = _rand((2, 8, 128), dtype, 11).cuda()
    kc = _rand((2, 3000, 2, 128), dtype, 12).cuda()
    vc = _rand((2, 3000, 2, 128), dtype, 13).cuda()
    sl = torch.tensor([3000, 1777], dtype=torch.int32).cuda()
    o1, l1 = sfa.decode(q, kc, vc, sl, num_splits=1)
    for s in (2, 7, 16, 0):
        o, l = sfa.decode(q, kc, vc, sl, num_splits=s)
        assert float((o.float() - o1.float()).abs().max()) <= 1e-3
        assert float((l - l1).abs().max()) <= 1e-4


def _paged_case(dtype, B, Hkv, D, page, blocks, seed):
    """A pool with more pages than the sequences use, and a block table that scatters every sequence over it."""
    g = torch.Generator().manual_seed(seed)
    pages = B * blocks + 5
    k_pool = _rand((pages, page, Hkv, D), dtype, seed + 1)
    v_pool = _rand((pages, page, Hkv, D), dtype, seed + 2)
    table = torch.randperm(pages, generator=g)[:B * blocks].reshape(B, blocks).to(torch.int32)
    return k_pool, v_pool, table


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("B,Hq,Hkv,D,page,blocks,lens,splits", [
    (2, 8, 2, 128, 16, 40, [640, 333], 0),       # smallest page: one 16-key step per page
    (3, 8, 1, 128, 64, 11, [1, 64, 700], 5),     # GQA 8, one-key sequence, page boundary, odd split count
    (2, 4, 2, 64, 128, 5, [513, 17], 3),         # head_dim 64
    (2, 2, 2, 128, 256, 2, [0, 512], 4),         # an empty sequence; chunk (64) smaller than the page
    (1, 32, 8, 128, 256, 16, [4096], 0),         # the C4 head layout
    (2, 16, 1, 128, 32, 20, [640, 35], 4),       # MQA 16 (two chunks of 8 query heads) over pages
])
def test_decode_paged_matches_oracle_and_dense_kernel(dtype, B, Hq, Hkv, D, page, blocks, lens, splits):
    """Paged cache (scope row f1): against the oracle over the gathered cache, and bit for bit against the dense
    kernel with the same split count (the page table changes addresses, never arithmetic)."""
    import starflashattention_b200 as sfa
    q = _rand((B, Hq, D), dtype, 31)
    k_pool, v_pool, table = _paged_case(dtype, B, Hkv, D, page, blocks, 32)
    seqlens = torch.tensor(lens, dtype=torch.int32)
    kc, vc = naive.gather_pages(k_pool, table), naive.gather_pages(v_pool, table)
    o_ref, lse_ref = naive.decode_attention(q, kc, vc, seqlens)
    o, lse = sfa.decode_paged(q.cuda(), k_pool.cuda(), v_pool.cuda(), table.cuda(), seqlens.cuda(), num_splits=splits)
    torch.cuda.synchronize()
    _assert_close(o, o_ref, dtype, "O")
    fin = torch.isfinite(lse_ref)
    assert torch.equal(torch.isfinite(lse.cpu()), fin)
    assert float((lse.cpu()[fin] - lse_ref[fin]).abs().max() if fin.any() else 0.0) <= 1e-3
    s = splits or sfa.decode_num_splits(B, blocks * page, Hkv)
    o_d, lse_d = sfa.decode(q.cuda(), kc.cuda(), vc.cuda(), seqlens.cuda(), num_splits=s)
    o_p, lse_p = sfa.decode_paged(q.cuda(), k_pool.cuda(), v_pool.cuda(), table.cuda(), seqlens.cuda(), num_splits=s)
    assert torch.equal(o_p, o_d) and torch.equal(lse_p, lse_d)


def test_decode_paged_shared_pages_and_strided_table():
    """Two sequences may name the same pool pages (a shared prefix), and the table may be a view with a row stride."""
    import starflashattention_b200 as sfa
    dtype = torch.bfloat16
    q = _rand((2, 8, 128), dtype, 41)
    k_pool, v_pool, _ = _paged_case(dtype, 2, 2, 128, 64, 8, 42)
    wide = torch.zeros((2, 12), dtype=torch.int32)
    wide[0, :8] = torch.tensor([3, 1, 4, 15, 9, 2, 6, 5])
    wide[1, :8] = torch.tensor([3, 1, 4, 15, 20, 19, 18, 17])      # first four pages shared
    table = wide[:, :8]                                           # row stride 12
    seqlens = torch.tensor([500, 301], dtype=torch.int32)
    o_ref, lse_ref = naive.decode_attention(q, naive.gather_pages(k_pool, table), naive.gather_pages(v_pool, table), seqlens)
    o, lse = sfa.decode_paged(q.cuda(), k_pool.cuda(), v_pool.cuda(), wide.cuda()[:, :8], seqlens.cuda())
    _assert_close(o, o_ref, dtype, "O")
    assert float((lse.cpu() - lse_ref).abs().max()) <= 1e-3


def test_decode_paged_rejects_bad_arguments():
    import starflashattention_b200 as sfa
    from starflashattention_b200._capi import StarFlashAttentionError
    q = torch.zeros((1, 4, 128), dtype=torch.float16, device="cuda")
    sl = torch.tensor([10], dtype=torch.int32, device="cuda")
    table = torch.zeros((1, 2), dtype=torch.int32, device="cuda")
    for page in (8, 24):                                          # below 16, not a power of two
        pool = torch.zeros((4, page, 4, 128), dtype=torch.float16, device="cuda")
        with pytest.raises(StarFlashAttentionError):
            sfa.decode_paged(q, pool, pool, table, sl)
    pool = torch.zeros((4, 16, 4, 128), dtype=torch.float16, device="cuda")
    with pytest.raises(ValueError):
        sfa.decode_paged(q, pool, pool, table.to(torch.int64), sl)
    with pytest.raises(ValueError):
        sfa.decode_paged(q, pool, pool, table, None)


def test_decode_paged_full_c4_shape_equals_dense():
    """BASELINE config 4 at full size through a shuffled page table (page 256): bit-identical to the dense kernel."""
    import starflashattention_b200 as sfa
    dtype, B, Hq, Hkv, D, M, page = torch.bfloat16, 8, 32, 8, 128, 32768, 256
    g = torch.Generator(device="cuda").manual_seed(5)
    blocks = M // page
    k_pool = torch.randn((B * blocks, page, Hkv, D), generator=g, device="cuda", dtype=torch.float32).to(dtype)
    v_pool = torch.randn((B * blocks, page, Hkv, D), generator=g, device="cuda", dtype=torch.float32).to(dtype)
    q = torch.randn((B, Hq, D), generator=g, device="cuda", dtype=torch.float32).to(dtype)
    table = torch.randperm(B * blocks, generator=g, device="cuda").reshape(B, blocks).to(torch.int32)
    sl = torch.tensor([M, M - 1, M - 255, 16384, 20000, 31111, M, 257], dtype=torch.int32, device="cuda")
    o_p, lse_p = sfa.decode_paged(q, k_pool, v_pool, table, sl)
    kc = k_pool[table.long().reshape(-1)].reshape(B, M, Hkv, D)
    vc = v_pool[table.long().reshape(-1)].reshape(B, M, Hkv, D)
    o_d, lse_d = sfa.decode(q, kc, vc, sl)
    torch.cuda.synchronize()
    assert torch.isfinite(o_p.float()).all() and torch.isfinite(lse_p).all()
    assert torch.equal(o_p, o_d) and torch.equal(lse_p, lse_d)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("Hq,Hkv,D,page,use_tables", [(8, 2, 128, 64, True), (4, 4, 128, 16, False), (8, 1, 64, 128, True),
                                                     (32, 2, 128, 64, True)])   # GQA 16: both chunks append the same row
def test_decoder_step_paged_matches_oracle_and_appends_into_the_page(dtype, Hq, Hkv, D, page, use_tables):
    """Fused RoPE + append + attention over a paged cache with GQA: positions at a page start, a page end, mid-page
    and on an empty sequence; the appended rows land in the page the block table names and nothing else changes."""
    import starflashattention_b200 as sfa
    B, blocks = 4, 6
    max_len = blocks * page
    k_pool, v_pool, table = _paged_case(dtype, B, Hkv, D, page, blocks, 51)
    qkv = _rand((B, Hq + 2 * Hkv, D), dtype, 55)
    seq = torch.tensor([2 * page, 3 * page - 1, page + 7, 0], dtype=torch.int32)
    kc, vc = naive.gather_pages(k_pool, table), naive.gather_pages(v_pool, table)
    q, k, v = qkv[:, :Hq], qkv[:, Hq:Hq + Hkv], qkv[:, Hq + Hkv:]
    o_ref, lse_ref, k_rot = naive.decode_step_gqa(q, k, v, kc, vc, seq, table_dtype=dtype if use_tables else None)
    cos = sin = None
    if use_tables:
        cos, sin = sfa.compute_rotary_table(max_len, D, dtype=dtype)
    kp, vp = k_pool.cuda(), v_pool.cuda()
    lse = torch.empty((B, Hq), dtype=torch.float32, device="cuda")
    o = sfa.decoder_step_paged(qkv.cuda(), kp, vp, table.cuda(), seq.cuda(), Hq, rotary_cos_table=cos,
                               rotary_sin_table=sin, lse=lse, num_splits=3)
    torch.cuda.synchronize()
    _assert_close(o, o_ref, dtype, "O")
    assert float((lse.cpu() - lse_ref).abs().max()) <= 1e-3
    k_want, v_want = k_pool.clone(), v_pool.clone()
    for b in range(B):
        pos = int(seq[b])
        pg, row = int(table[b, pos // page]), pos % page
        got_k = kp[pg, row].cpu()
        assert float((got_k.float() - k_rot[b].float()).abs().max()) <= 1e-2
        k_want[pg, row] = got_k
        v_want[pg, row] = v[b]
    assert torch.equal(kp.cpu(), k_want) and torch.equal(vp.cpu(), v_want)
    # the next token attends over the row just appended: the dense decode over the gathered, updated cache agrees
    q2 = _rand((B, Hq, D), dtype, 56).cuda()
    o_p, lse_p = sfa.decode_paged(q2, kp, vp, table.cuda(), (seq + 1).cuda(), num_splits=2)
    o_d, lse_d = sfa.decode(q2, naive.gather_pages(kp.cpu(), table).cuda(), naive.gather_pages(vp.cpu(), table).cuda(),
                            (seq + 1).cuda(), num_splits=2)
    assert torch.equal(o_p, o_d) and torch.equal(lse_p, lse_d)


def test_decoder_step_paged_equals_dense_drop_in_step():
    """MHA, one layer: the paged fused step and `mha_fwd_cuda` over the gathered dense cache give the same bits."""
    import starflashattention_b200 as sfa
    dtype, B, H, D, page, blocks = torch.float16, 2, 4, 128, 64, 8
    M = page * blocks
    k_pool, v_pool, table = _paged_case(dtype, B, H, D, page, blocks, 61)
    qkv = _rand((B, 3, H, D), dtype, 65)
    seq = torch.tensor([300, 511], dtype=torch.int32)
    kc = naive.gather_pages(k_pool, table)[:, None].contiguous().cuda()      # (B, L=1, M, H, D)
    vc = naive.gather_pages(v_pool, table)[:, None].contiguous().cuda()
    o_d = torch.empty((B, H, D), dtype=dtype, device="cuda")
    sfa.mha_fwd_cuda(qkv.cuda(), None, None, None, kc, vc, seq.cuda(), o_d, B, M, H, D, D, M, 1, 0, num_splits=4)
    kp, vp = k_pool.cuda(), v_pool.cuda()
    o_p = sfa.decoder_step_paged(qkv.reshape(B, 3 * H, D).cuda(), kp, vp, table.cuda(), seq.cuda(), H, num_splits=4)
    torch.cuda.synchronize()
    assert torch.equal(o_p, o_d)
    assert torch.equal(naive.gather_pages(kp.cpu(), table), kc[:, 0].cpu())
    assert torch.equal(naive.gather_pages(vp.cpu(), table), vc[:, 0].cpu())


def test_decoder_step_paged_full_cache_does_not_append():
    import starflashattention_b200 as sfa
    dtype, B, Hq, Hkv, D, page, blocks = torch.bfloat16, 1, 8, 2, 128, 32, 4
    k_pool, v_pool, table = _paged_case(dtype, B, Hkv, D, page, blocks, 71)
    qkv = _rand((B, Hq + 2 * Hkv, D), dtype, 75)
    seq = torch.tensor([page * blocks], dtype=torch.int32)                   # no room for the new token
    kp, vp = k_pool.cuda(), v_pool.cuda()
    o = sfa.decoder_step_paged(qkv.cuda(), kp, vp, table.cuda(), seq.cuda(), Hq)
    torch.cuda.synchronize()
    assert torch.equal(kp.cpu(), k_pool) and torch.equal(vp.cpu(), v_pool)
    assert bool(torch.isfinite(o.float()).all())


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("B,Hq,Hkv,D,M,lens", [
    (2, 8, 2, 128, 3000, [3000, 1777]),
    (3, 8, 1, 64, 700, [1, 64, 700]),            # splits without keys take part in the cluster merge with weight 0
    (2, 4, 4, 128, 512, [0, 512]),               # an empty sequence: O = 0, LSE = -inf
])
def test_decode_cluster_merge_matches_two_launch_form(dtype, B, Hq, Hkv, D, M, lens, tuning):
    """2 .. 8 splits merge inside a thread-block cluster through distributed shared memory (one launch); the
    two-launch form (partials in global memory + combine kernel, SFA_DECODE_CLUSTER=0) is the same arithmetic up
    to the order of the fp32 sums.  Both against the oracle."""
    import starflashattention_b200 as sfa
    q, kc, vc = _rand((B, Hq, D), dtype, 91), _rand((B, M, Hkv, D), dtype, 92), _rand((B, M, Hkv, D), dtype, 93)
    sl = torch.tensor(lens, dtype=torch.int32)
    o_ref, lse_ref = naive.decode_attention(q, kc, vc, sl)
    fin = torch.isfinite(lse_ref)
    qd, kd, vd, sd = q.cuda(), kc.cuda(), vc.cuda(), sl.cuda()
    for s in (2, 3, 5, 8):
        res = {}
        for cl in ("1", "0"):
            tuning("decode_cluster", int(cl), 1)
            n0 = sfa.launch_count()
            o, lse = sfa.decode(qd, kd, vd, sd, num_splits=s)
            torch.cuda.synchronize()
            res[cl] = (o.float().cpu(), lse.cpu(), sfa.launch_count() - n0)
        assert res["1"][2] == 1 and res["0"][2] == 2
        _assert_close(res["1"][0], o_ref, dtype, f"O, {s} cluster splits")
        assert torch.equal(torch.isfinite(res["1"][1]), fin)
        assert float((res["1"][1][fin] - lse_ref[fin]).abs().max()) <= 1e-3
        assert float((res["1"][0] - res["0"][0]).abs().max()) <= 1e-3
        assert float((res["1"][1][fin] - res["0"][1][fin]).abs().max()) <= 1e-5
    tuning("decode_cluster", 1, 1)
    n0 = sfa.launch_count()
    sfa.decode(qd, kd, vd, sd, num_splits=9)      # more than a portable cluster: two launches
    assert sfa.launch_count() - n0 == 2


def test_decode_outlier_score_does_not_overflow():
    import starflashattention_b200 as sfa
    dtype = torch.float16
    q = _rand((1, 4, 128), dtype, 21)
    kc = _rand((1, 600, 4, 128), dtype, 22)
    vc = _rand((1, 600, 4, 128), dtype, 23)
    kc[0, 311, 2] = q[0, 2] * 3.0          # a score of ~ +30 after scaling
    sl = torch.tensor([600], dtype=torch.int32)
    o_ref, lse_ref = naive.decode_attention(q, kc, vc, sl)
    o, lse = sfa.decode(q.cuda(), kc.cuda(), vc.cuda(), sl.cuda())
    assert torch.isfinite(o).all() and torch.isfinite(lse).all()
    _assert_close(o, o_ref, dtype, "O")
    assert float((lse.cpu() - lse_ref).abs().max()) <= 1e-3


def test_decode_full_c4_shape_spot_check():
    """BASELINE config 4 at full size (B=8, 32 q-heads / 8 kv-heads, KV 32768, bf16): two sequences are
    checked against the oracle, all of them for finiteness and split-invariance of the LSE."""
    import starflashattention_b200 as sfa
    dtype = torch.bfloat16
    B, Hq, Hkv, D, M = 8, 32, 8, 128, 32768
    g = torch.Generator(device="cuda").manual_seed(4)
    kc = torch.randn((B, M, Hkv, D), generator=g, device="cuda", dtype=torch.float32).to(dtype)
    vc = torch.randn((B, M, Hkv, D), generator=g, device="cuda", dtype=torch.float32).to(dtype)
    q = torch.randn((B, Hq, D), generator=g, device="cuda", dtype=torch.float32).to(dtype)
    lens = torch.tensor([32768, 16385, 32768, 20000, 32768, 32768, 31999, 32768], dtype=torch.int32)
    o, lse = sfa.decode(q, kc, vc, lens.cuda())
    o1, lse1 = sfa.decode(q, kc, vc, lens.cuda(), num_splits=3)
    torch.cuda.synchronize()
    assert torch.isfinite(o).all() and torch.isfinite(lse).all()
    assert float((lse - lse1).abs().max()) <= 1e-4
    for b in (1, 6):
        o_ref, lse_ref = naive.decode_attention(q[b:b + 1].cpu(), kc[b:b + 1].cpu(), vc[b:b + 1].cpu(), lens[b:b + 1])
        _assert_close(o[b:b + 1], o_ref, dtype, f"O[{b}]")
        assert float((lse[b:b + 1].cpu() - lse_ref).abs().max()) <= 1e-3


# ----------------------------------------------------------------------------- drop-in decoder step
def _step_bound(qkv, kc, vc, seq, layer, dtype, table_dtype=None):
    """Error bound of the decoder step against the oracle, per output element.  fp16: north_star's 2e-3.  bf16: plus
    the rounding of P (tensor-core operand) and of O (half an ulp = 2^-9 relative each) against the oracle's
    P @ |V| on the same inputs (the prefill bound of tests/test_fwd_gpu.py).  Rotated q and k are rounded to the data
    type by kernel and oracle alike; where the two fp32 rotations differ in the last bit an element of q or k may round
    to the neighbouring value (one ulp = 2^-7 relative in bf16, 2^-10 in fp16), which moves a score by at most
    ulp * max|q_i k_i| * scale and P (hence O) by that fraction of P @ |V|: the third term, for up to 8 such elements."""
    qa, va = qkv.clone(), vc.abs()
    qa[:, 2] = qa[:, 2].abs()
    mag, *_ = naive.decode_step_reference(qa, kc, va, seq, layer, table_dtype=table_dtype)
    mag = mag.float()
    ulp = 2.0 ** -10 if dtype == torch.float16 else 2.0 ** -7
    D = qkv.shape[-1]
    flip = 8 * ulp * float(qkv[:, 0].float().abs().max()) * float(max(qkv[:, 1].float().abs().max(), kc.float().abs().max())) / D ** 0.5
    base = 2e-3 if dtype == torch.float16 else 2e-3 + 2.0 ** -8 * mag
    return base + flip * mag


def _load_golden(path):
    z = np.load(path, allow_pickle=False)
    B, H, D, L, M, cur, layer = [int(x) for x in z["meta"]]
    dt = DT[str(z["dtype"])]
    return z, (B, H, D, L, M, cur, layer), dt


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "ref_decode_*.npz"))),
                         ids=lambda p: os.path.basename(p)[11:-4])
def test_decoder_step_matches_reference_fixtures(path):
    """mha_fwd_cuda against outputs of the reference's own LlamaAttention.forward (tools/make_golden.py)."""
    import starflashattention_b200 as sfa
    z, (B, H, D, L, M, cur, layer), dt = _load_golden(path)
    run_dt = torch.float16 if dt == torch.float32 else dt     # the library computes in fp16/bf16 only
    if str(z["fill"]) == "ones":
        qkv = torch.ones(B, 3, H, D, dtype=run_dt)
        kc = torch.ones(B, L, M, H, D, dtype=run_dt)
        vc = torch.ones(B, L, M, H, D, dtype=run_dt)
    else:
        qkv = torch.from_numpy(z["qkv"]).to(run_dt)
        kc = torch.zeros(B, L, M, H, D, dtype=run_dt)
        vc = torch.zeros(B, L, M, H, D, dtype=run_dt)
        kc[:, layer, :cur - 1] = torch.from_numpy(z["k_cache_layer"]).to(run_dt)
        vc[:, layer, :cur - 1] = torch.from_numpy(z["v_cache_layer"]).to(run_dt)
    seq = torch.full((B,), cur - 1, dtype=torch.int32)
    o = torch.empty(B, H, D, dtype=run_dt, device="cuda")
    kc_d, vc_d = kc.cuda(), vc.cuda()
    ret = sfa.mha_fwd_cuda(qkv.cuda(), None, None, None, kc_d, vc_d, seq.cuda(), o, B, M, H, D, D, M, L, layer)
    torch.cuda.synchronize()
    assert ret.data_ptr() == o.data_ptr()                      # written in place and returned (flash_api.cpp:67)
    ref = torch.from_numpy(z["o_ref_fp32"])[:, :, 0, :]
    if dt == torch.float32:
        # fixture inputs are fp32; the library rounds them to fp16 first, so compare with the oracle on rounded inputs
        ref, *_ = naive.decode_step_reference(qkv, kc, vc, seq, layer)
    err = (o.float().cpu() - ref.float()).abs()
    assert bool((err <= _step_bound(qkv, kc, vc, seq, layer, run_dt)).all()), float(err.max())


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("use_tables", [False, True])
def test_decoder_step_ragged_positions_and_cache_append(dtype, use_tables):
    import starflashattention_b200 as sfa
    B, L, M, H, D, layer = 3, 2, 640, 8, 128, 1
    qkv = _rand((B, 3, H, D), dtype, 31)
    kc = _rand((B, L, M, H, D), dtype, 32)
    vc = _rand((B, L, M, H, D), dtype, 33)
    seq = torch.tensor([0, 255, 639], dtype=torch.int32)
    cos = sin = None
    if use_tables:
        cos, sin = sfa.compute_rotary_table(M, D, dtype=dtype)
    o_ref, lse_ref, k_new, v_new = naive.decode_step_reference(
        qkv, kc, vc, seq, layer, table_dtype=dtype if use_tables else None)
    kc_d, vc_d = kc.cuda(), vc.cuda()
    o = torch.empty(B, H, D, dtype=dtype, device="cuda")
    lse = torch.empty(B, H, dtype=torch.float32, device="cuda")
    sfa.mha_fwd_cuda(qkv.cuda(), None, None, None, kc_d, vc_d, seq.cuda(), o, B, M, H, D, D, M, L, layer,
                     rotary_cos_table=cos, rotary_sin_table=sin, lse=lse)
    torch.cuda.synchronize()
    err = (o.float().cpu() - o_ref.float()).abs()
    assert bool((err <= _step_bound(qkv, kc, vc, seq, layer, dtype, table_dtype=dtype if use_tables else None)).all()), float(err.max())
    assert float((lse.cpu() - lse_ref).abs().max()) <= 2e-3
    ulp = 2.0 ** -10 if dtype == torch.float16 else 2.0 ** -7
    for b in range(B):
        pos = int(seq[b])
        got_k = kc_d[b, layer, pos].float().cpu()
        # rotated key: one rounding of the same fp32 expression; angles at large positions differ by fp32 ulps
        assert float((got_k - k_new[b].float()).abs().max()) <= 4 * ulp * max(1.0, float(k_new[b].abs().max()))
        assert torch.equal(vc_d[b, layer, pos].cpu(), v_new[b])             # V is copied bit-exactly
        # nothing else in the cache moved
        other = torch.ones(M, dtype=torch.bool)
        other[pos] = False
        assert torch.equal(kc_d[b, layer][other].cpu(), kc[b, layer][other])
        assert torch.equal(kc_d[b, 1 - layer].cpu(), kc[b, 1 - layer])


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("D", [64, 128])
def test_decoder_step_fused_equals_three_launch_form(dtype, D, tuning):
    """RoPE + append inside the split-KV kernel (default) against the stand-alone RoPE/append kernel followed by
    the plain decode kernel (SFA_DECODER_FUSED=0): same arithmetic, so outputs and caches must be bit-identical.
    Positions cover an empty cache, a step boundary (16 keys per warp step), a chunk boundary and a full cache."""
    import starflashattention_b200 as sfa
    B, L, M, H, layer = 6, 2, 1024, 4, 0
    qkv = _rand((B, 3, H, D), dtype, 41)
    kc = _rand((B, L, M, H, D), dtype, 42)
    vc = _rand((B, L, M, H, D), dtype, 43)
    seq = torch.tensor([0, 15, 16, 511, 777, 1023], dtype=torch.int32)
    cos, sin = sfa.compute_rotary_table(M, D, dtype=dtype)
    outs = []
    for fused in ("1", "0"):
        tuning("decoder_fused", int(fused), 1)
        kc_d, vc_d = kc.cuda(), vc.cuda()
        o = torch.empty(B, H, D, dtype=dtype, device="cuda")
        lse = torch.empty(B, H, dtype=torch.float32, device="cuda")
        n0 = sfa.launch_count()
        sfa.mha_fwd_cuda(qkv.cuda(), None, None, None, kc_d, vc_d, seq.cuda(), o, B, M, H, D, D, M, L, layer,
                         rotary_cos_table=cos, rotary_sin_table=sin, lse=lse, num_splits=4)
        torch.cuda.synchronize()
        outs.append((o.cpu(), lse.cpu(), kc_d.cpu(), vc_d.cpu(), sfa.launch_count() - n0))
    (o1, l1, k1, v1, n1), (o0, l0, k0, v0, n0_) = outs
    # fused: one split-KV launch whose 4 splits merge in their cluster; unfused: RoPE/append kernel first
    assert n1 == 1 and n0_ == 2
    assert torch.equal(k1, k0) and torch.equal(v1, v0)
    assert torch.equal(o1, o0) and torch.equal(l1, l0)


def test_decoder_step_cache_full_does_not_append():
    """seq_len == memory_max_len: no room for the token - nothing is written, attention runs over the full cache."""
    import starflashattention_b200 as sfa
    B, L, M, H, D = 1, 1, 256, 2, 128
    qkv = _rand((B, 3, H, D), torch.float16, 51)
    kc = _rand((B, L, M, H, D), torch.float16, 52)
    vc = _rand((B, L, M, H, D), torch.float16, 53)
    kc_d, vc_d = kc.cuda(), vc.cuda()
    o = torch.empty(B, H, D, dtype=torch.float16, device="cuda")
    sfa.mha_fwd_cuda(qkv.cuda(), None, None, None, kc_d, vc_d, torch.tensor([M], dtype=torch.int32).cuda(), o,
                     B, M, H, D, D, M, L, 0)
    torch.cuda.synchronize()
    assert torch.equal(kc_d.cpu(), kc) and torch.equal(vc_d.cpu(), vc)
    assert bool(torch.isfinite(o.float()).all())


def test_decoder_step_graph_replay_matches_eager_steps():
    """Eight tokens through a two-layer captured decoder step (one graph launch per token) against the same
    tokens through eager calls: identical outputs and caches; the sequence lengths advance on the device."""
    import starflashattention_b200 as sfa
    dtype = torch.bfloat16
    B, L, M, H, D, steps = 2, 2, 512, 8, 128, 8
    kc = _rand((B, L, M, H, D), dtype, 61)
    vc = _rand((B, L, M, H, D), dtype, 62)
    seq0 = torch.tensor([100, 37], dtype=torch.int32)
    toks = [_rand((L, B, 3, H, D), dtype, 70 + i).cuda() for i in range(steps)]
    # eager
    kc_e, vc_e, seq_e = kc.cuda(), vc.cuda(), seq0.cuda()
    outs_e = []
    for t in toks:
        o = torch.empty(L, B, H, D, dtype=dtype, device="cuda")
        for layer in range(L):
            sfa.mha_fwd_cuda(t[layer], None, None, None, kc_e, vc_e, seq_e, o[layer], B, M, H, D, D, M, L, layer)
        seq_e += 1
        outs_e.append(o.clone())
    # captured
    kc_g, vc_g = kc.cuda(), vc.cuda()
    g = sfa.DecoderStepGraph(kc_g, vc_g, seq0.cuda())
    assert torch.equal(kc_g.cpu(), kc) and torch.equal(vc_g.cpu(), vc)      # capture left the caches untouched
    n0 = sfa.launch_count()
    for t, o_e in zip(toks, outs_e):
        o = g.step(t)
        torch.cuda.synchronize()
        assert torch.equal(o.cpu(), o_e.cpu())
    assert sfa.launch_count() == n0            # replays launch through the graph, not through the library's host code
    assert torch.equal(g.seq_len.cpu(), seq0 + steps)
    assert torch.equal(kc_g.cpu(), kc_e.cpu()) and torch.equal(vc_g.cpu(), vc_e.cpu())


def test_rotary_table_and_fill():
    import starflashattention_b200 as sfa
    M, D = 2048, 128
    cos, sin = sfa.compute_rotary_table(M, D, dtype=torch.float16)
    pos = torch.arange(M, dtype=torch.float64)[:, None]
    inv = 1.0 / (10000.0 ** (torch.arange(0, D, 2, dtype=torch.float64) / D))[None]
    ang = pos * inv
    # fp32 angle rounding grows with the position: |d angle| <= pos * 2^-23
    tol = 2.0 ** -10 + pos * 2.0 ** -22
    assert bool(((cos.double().cpu() - ang.cos()).abs() <= tol).all())
    assert bool(((sin.double().cpu() - ang.sin()).abs() <= tol).all())
    a = torch.zeros(100003, dtype=torch.float16, device="cuda")
    sfa.init_half_array(a, 1.0)
    assert bool((a == 1.0).all())
    b = torch.zeros(77, dtype=torch.bfloat16, device="cuda")
    sfa.init_half_array(b, -2.5)
    assert bool((b == -2.5).all())


def test_all_ones_fixture_of_the_cpp_example():
    """examples/cpp/testFlashDecoder.cc:63-78,138-150: all-ones inputs, B=2 H=32 D=128, 4 layers, layer 0."""
    import starflashattention_b200 as sfa
    B, H, D, L = 2, 32, 128, 4
    for M, n in [(512, 511), (1024, 1023), (2048, 2047), (8192, 6143), (8192, 8191)]:
        qkv = torch.ones(B, 3, H, D, dtype=torch.float16, device="cuda")
        kc = torch.ones(B, L, M, H, D, dtype=torch.float16, device="cuda")
        vc = torch.ones(B, L, M, H, D, dtype=torch.float16, device="cuda")
        seq = torch.full((B,), n, dtype=torch.int32, device="cuda")
        o = torch.zeros(B, H, D, dtype=torch.float16, device="cuda")
        cos, sin = sfa.compute_rotary_table(M, D)
        sfa.mha_fwd_cuda(qkv, None, None, None, kc, vc, seq, o, B, M, H, D, D, M, L, 0,
                         rotary_cos_table=cos, rotary_sin_table=sin)
        assert float((o.float() - 1.0).abs().max()) <= 1e-3


def test_torch_extension_decoder_step_matches_oracle(torch_ext):
    """The pybind entry the reference's Python example calls (examples/python/testFlashDecoder.py:124-126)."""
    B, L, M, H, D = 2, 3, 640, 8, 128
    g = torch.Generator().manual_seed(77)
    qkv = torch.randn(B, 3, H, D, generator=g).half()
    kt = torch.randn(B, L, M, H, D, generator=g).half()
    vt = torch.randn(B, L, M, H, D, generator=g).half()
    seq = torch.tensor([511, 130], dtype=torch.int32)
    o_ref, _, _, _ = naive.decode_step_reference(qkv, kt, vt, seq, 2)
    o = torch.empty(B, H, D, dtype=torch.float16, device="cuda")
    bias = torch.empty(0, device="cuda")
    n0 = torch_ext.launch_count()
    ret = torch_ext.mha_fwd_cuda(qkv=qkv.cuda(), q_bias=bias, k_bias=bias, v_bias=bias, k_cache_table=kt.cuda(),
                                 v_cache_table=vt.cuda(), seq_len=seq.cuda(), o=o, batch_size=B, memory_max_len=M,
                                 num_heads=H, head_dim=D, rotary_embedding_dim=D, max_input_length=M, num_layer=L,
                                 idx_layer=2)
    torch.cuda.synchronize()
    assert ret.data_ptr() == o.data_ptr()
    assert torch_ext.launch_count() > n0
    assert float((o.float().cpu() - o_ref.float()).abs().max()) <= 3e-3


def test_torch_extension_fwd_and_decode(torch_ext):
    g = torch.Generator().manual_seed(78)
    q = torch.randn(1, 300, 4, 128, generator=g).half()
    k = torch.randn(1, 300, 2, 128, generator=g).half()
    v = torch.randn(1, 300, 2, 128, generator=g).half()
    o_ref, lse_ref = naive.attention_fwd(q, k, v, causal=True)
    o, lse = torch_ext.fwd(q.cuda(), k.cuda(), v.cuda(), causal=True)
    assert float((o.float().cpu() - o_ref.float()).abs().max()) <= 2e-3
    assert float((lse.cpu() - lse_ref).abs().max()) <= 1e-3
    qd = torch.randn(2, 8, 128, generator=g).half()
    kc = torch.randn(2, 1000, 2, 128, generator=g).half()
    vc = torch.randn(2, 1000, 2, 128, generator=g).half()
    sl = torch.tensor([1000, 17], dtype=torch.int32)
    o_ref, lse_ref = naive.decode_attention(qd, kc, vc, sl)
    o, lse = torch_ext.decode(qd.cuda(), kc.cuda(), vc.cuda(), sl.cuda())
    assert float((o.float().cpu() - o_ref.float()).abs().max()) <= 2e-3
    assert float((lse.cpu() - lse_ref).abs().max()) <= 1e-3
    # the same cache cut into 8 pages of 128 keys per sequence (1000 -> 1024 rows), pages in reverse order
    pad = torch.zeros(2, 24, 2, 128, dtype=torch.float16)
    k_pool = torch.cat([kc, pad], 1).reshape(16, 128, 2, 128).flip(0).contiguous()
    v_pool = torch.cat([vc, pad], 1).reshape(16, 128, 2, 128).flip(0).contiguous()
    table = (15 - torch.arange(16)).reshape(2, 8).to(torch.int32)
    o, lse = torch_ext.decode_paged(qd.cuda(), k_pool.cuda(), v_pool.cuda(), table.cuda(), sl.cuda())
    assert float((o.float().cpu() - o_ref.float()).abs().max()) <= 2e-3
    assert float((lse.cpu() - lse_ref).abs().max()) <= 1e-3


def test_decode_host_pipeline_matches_device_path():
    import starflashattention_b200 as sfa
    dtype = torch.bfloat16
    B, Hq, Hkv, D, M = 5, 8, 2, 128, 3000
    q = _rand((B, Hq, D), dtype, 41).pin_memory()
    kc = _rand((B, M, Hkv, D), dtype, 42).pin_memory()
    vc = _rand((B, M, Hkv, D), dtype, 43).pin_memory()
    sl = torch.tensor([3000, 1, 777, 2048, 129], dtype=torch.int32).pin_memory()
    o_h, lse_h = sfa.decode_host(q, kc, vc, sl, chunks=3)
    o_d, lse_d = sfa.decode(q.cuda(), kc.cuda(), vc.cuda(), sl.cuda())
    torch.cuda.synchronize()
    o_ref, lse_ref = naive.decode_attention(q, kc, vc, sl)
    _assert_close(o_h, o_ref, dtype, "decode_host O")
    assert float((lse_h - lse_ref).abs().max()) <= 1e-3
    # the split count may differ between a 2-sequence chunk and the 5-sequence batch: compare within rounding
    assert float((o_h.float() - o_d.float().cpu()).abs().max()) <= 2e-2
