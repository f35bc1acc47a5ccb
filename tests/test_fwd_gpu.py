"""GPU parity: the tcgen05/TMA prefill forward, through the C ABI, against the CPU oracle
(softmax(QK^T/sqrt(d))V of the reference's Python example, examples/python/testFlashDecoder.py:84-92).

Tolerances (north_star): |O - O_oracle| <= 2e-3 (fp16), LSE within 1e-3.  The oracle runs in fp32 on the same
fp16/bf16-rounded inputs.  bf16 carries 8 significand bits, so storing P (the tensor-core A operand) and O in
bf16 costs up to 2^-9 relative each: the bf16 bound is 2e-3 + 2^-8 * (P @ |V|), where P @ |V| is evaluated by
the oracle on the same inputs (an input-dependent, rigorous rounding bound; no slack beyond the 2e-3).
"""
import pytest
import torch

from oracle import naive

pytestmark = pytest.mark.gpu


def _rand(shape, dtype, seed, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(shape, generator=g) * scale).to(dtype)


def _oracle(q, k, v, **kw):
    """(O, LSE, P @ |V|) from the CPU oracle."""
    o_ref, lse_ref = naive.attention_fwd(q, k, v, **kw)
    mag, _ = naive.attention_fwd(q, k, v.abs(), **kw)
    return o_ref, lse_ref, mag


def _check(o, lse, o_ref, lse_ref, dtype, what="", mag=None):
    err = (o.float().cpu() - o_ref.float()).abs()
    if dtype == torch.float16:
        bound = torch.full_like(err, 2e-3)
    else:
        assert mag is not None, "bf16 checks need the oracle's P @ |V|"
        bound = 2e-3 + 2.0 ** -8 * mag.float()
    assert bool((err <= bound).all()), f"{what} O: max err {float(err.max()):.3e}"
    lse = lse.cpu()
    fin = torch.isfinite(lse_ref)
    assert torch.equal(torch.isfinite(lse), fin), f"{what} LSE finiteness"
    if fin.any():
        assert float((lse[fin] - lse_ref[fin]).abs().max()) <= 1e-3, f"{what} LSE"


@pytest.fixture(autouse=True)
def _watchdog():
    import starflashattention_b200 as sfa
    yield
    assert sfa.fwd_watchdog(clear=True) == 0, "prefill pipeline watchdog fired (an mbarrier wait timed out)"


CASES = [
    # B, Sq, Sk, Hq, Hkv, D, causal
    (1, 128, 128, 1, 1, 128, False),      # one tile
    (1, 128, 128, 1, 1, 128, True),
    (1, 256, 256, 2, 2, 128, False),      # two query tiles in one CTA
    (2, 512, 512, 4, 2, 128, True),       # GQA, several key tiles
    (1, 200, 333, 2, 1, 128, False),      # ragged tails on both axes
    (1, 333, 333, 2, 2, 128, True),
    (1, 100, 419, 2, 2, 128, True),       # Sq < Sk, bottom-right aligned mask
    (1, 300, 140, 1, 1, 128, True),       # Sq > Sk: leading rows see no key (O = 0, LSE = -inf)
    (1, 1, 257, 4, 4, 128, False),        # single query row
    (2, 384, 384, 2, 2, 64, False),       # head_dim 64
    (1, 640, 640, 4, 1, 64, True),
    (1, 1024, 1024, 8, 8, 128, True),
]


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("B,Sq,Sk,Hq,Hkv,D,causal", CASES)
def test_fwd_matches_oracle(dtype, B, Sq, Sk, Hq, Hkv, D, causal):
    import starflashattention_b200 as sfa
    q = _rand((B, Sq, Hq, D), dtype, 1)
    k = _rand((B, Sk, Hkv, D), dtype, 2)
    v = _rand((B, Sk, Hkv, D), dtype, 3)
    o_ref, lse_ref, mag = _oracle(q, k, v, causal=causal)
    o, lse = sfa.fwd(q.cuda(), k.cuda(), v.cuda(), causal=causal)
    torch.cuda.synchronize()
    _check(o, lse, o_ref, lse_ref, dtype, mag=mag)


def test_fwd_c1_shape_against_fp32_reference():
    """BASELINE config 1 (B=1 H=8 N=512 d=64): the repo's naive fp32 check, run in fp16 on the GPU."""
    import starflashattention_b200 as sfa
    q, k, v = (_rand((1, 512, 8, 64), torch.float16, s) for s in (5, 6, 7))
    o_ref, lse_ref = naive.attention_fwd(q, k, v)
    o, lse = sfa.fwd(q.cuda(), k.cuda(), v.cuda())
    _check(o, lse, o_ref, lse_ref, torch.float16, "C1")


def test_fwd_strided_views_and_scale():
    """q/k/v sliced out of a packed (B, S, 3, H, D) tensor, non-default softmax scale."""
    import starflashattention_b200 as sfa
    dtype = torch.bfloat16
    qkv = _rand((2, 300, 3, 4, 128), dtype, 9)
    q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
    o_ref, lse_ref, mag = _oracle(q, k, v, causal=True, softmax_scale=0.05)
    d = qkv.cuda()
    o, lse = sfa.fwd(d[:, :, 0], d[:, :, 1], d[:, :, 2], causal=True, softmax_scale=0.05)
    _check(o, lse, o_ref, lse_ref, dtype, "strided", mag=mag)


def test_fwd_outlier_scores():
    """A +30 outlier late in the sequence forces the lazy O rescale; heavy-tailed inputs."""
    import starflashattention_b200 as sfa
    dtype = torch.float16
    q = _rand((1, 384, 2, 128), dtype, 13)
    k = _rand((1, 384, 2, 128), dtype, 14)
    v = _rand((1, 384, 2, 128), dtype, 15)
    k[0, 300, 0] = q[0, 17, 0] * 3.0
    k[0, 10, 1] = q[0, 200, 1] * 2.5
    o_ref, lse_ref = naive.attention_fwd(q, k, v)
    o, lse = sfa.fwd(q.cuda(), k.cuda(), v.cuda())
    assert torch.isfinite(o).all()
    _check(o, lse, o_ref, lse_ref, dtype, "outlier")


def test_fwd_causal_prefix_property():
    """Size-independent property: under the causal mask, row i depends on keys <= i only, so the first n
    rows of a long run equal a run on the n-row prefix."""
    import starflashattention_b200 as sfa
    dtype = torch.bfloat16
    q, k, v = (_rand((1, 2048, 2, 128), dtype, s).cuda() for s in (21, 22, 23))
    o_full, lse_full = sfa.fwd(q, k, v, causal=True)
    n = 768
    o_pre, lse_pre = sfa.fwd(q[:, :n], k[:, :n], v[:, :n], causal=True)
    # identical tiles in identical order: only the lazy-rescale decisions could differ, and they do not here
    assert float((o_full[:, :n].float() - o_pre.float()).abs().max()) <= 2e-3 + 2.0 ** -8 * 4.0
    assert float((lse_full[:, :, :n] - lse_pre).abs().max()) <= 1e-4


def _spot_check(q, k, v, o, lse, causal, rows, heads, dtype):
    """Oracle on sampled (batch, head) slices and query rows of a large problem."""
    g = q.shape[2] // k.shape[2]
    Sq, Sk = q.shape[1], k.shape[1]
    for (b, h) in heads:
        qs = q[b:b + 1, rows, h:h + 1].cpu()
        ks, vs = k[b:b + 1, :, h // g:h // g + 1].cpu(), v[b:b + 1, :, h // g:h // g + 1].cpu()
        for i, r in enumerate(rows):
            lim = r + (Sk - Sq) + 1 if causal else Sk
            o_ref, lse_ref, mag = _oracle(qs[:, i:i + 1], ks[:, :lim], vs[:, :lim])
            _check(o[b:b + 1, r:r + 1, h:h + 1], lse[b:b + 1, h:h + 1, r:r + 1], o_ref, lse_ref, dtype,
                   f"b{b} h{h} r{r}", mag=mag)


def _sample_rows(S, seed, n_random):
    """Fixed rows around every tile / block boundary the kernel has (128-row query tiles, 256-row CTA blocks:
    first and last rows, the SECOND query tile of the first, a middle and the last block) plus seeded random rows."""
    g = torch.Generator().manual_seed(seed)
    fixed = [0, 127, 128, 255, 256, S // 2 - 1, S // 2, S // 2 + 129, S - 256, S - 256 + 5, S - 128, S - 128 + 100, S - 1]
    rnd = torch.randint(0, S, (n_random,), generator=g).tolist()
    return sorted(set(fixed + rnd))


def test_fwd_full_c3_shape_spot_check():
    """BASELINE config 3 at full size (causal bf16 B=8 H=32 N=8192 d=128): 25 rows (tile boundaries + a seeded
    random dozen) of 8 (batch, head) pairs against the oracle, O and LSE; everything for finiteness."""
    import starflashattention_b200 as sfa
    dtype = torch.bfloat16
    B, S, H, D = 8, 8192, 32, 128
    g = torch.Generator(device="cuda").manual_seed(3)
    q, k, v = (torch.randn((B, S, H, D), generator=g, device="cuda", dtype=torch.float32).to(dtype) for _ in range(3))
    o, lse = sfa.fwd(q, k, v, causal=True)
    torch.cuda.synchronize()
    assert sfa.fwd_watchdog() == 0
    assert torch.isfinite(o).all() and torch.isfinite(lse).all()
    rows = _sample_rows(S, 33, 12)
    assert len(rows) >= 24
    heads = [(0, 0), (7, 31), (3, 17), (1, 5), (2, 30), (4, 11), (5, 24), (6, 8)]
    _spot_check(q, k, v, o, lse, True, rows, heads, dtype)


def test_fwd_full_c5_shape_spot_check():
    """BASELINE config 5 at full size on one GPU (causal bf16 B=1 H=32 N=65536 d=128: 256 query blocks per head, up
    to 512 key tiles per query tile, 256 wraps of the K/V ring per item, TMA row coordinates up to 65408): 25 rows
    (including rows of the SECOND query tile of late blocks: 65408+100, 65535, 32768+129) of 5 heads against the
    oracle, O and LSE; everything for finiteness; watchdog clean."""
    import starflashattention_b200 as sfa
    dtype = torch.bfloat16
    B, S, H, D = 1, 65536, 32, 128
    g = torch.Generator(device="cuda").manual_seed(5)
    q, k, v = (torch.randn((B, S, H, D), generator=g, device="cuda", dtype=torch.float32).to(dtype) for _ in range(3))
    o, lse = sfa.fwd(q, k, v, causal=True)
    torch.cuda.synchronize()
    assert sfa.fwd_watchdog() == 0
    assert torch.isfinite(o).all() and torch.isfinite(lse).all()
    rows = _sample_rows(S, 55, 12)
    assert len(rows) >= 24 and 65535 in rows and 65408 + 100 in rows and 32768 + 129 in rows
    _spot_check(q, k, v, o, lse, True, rows, [(0, 0), (0, 7), (0, 16), (0, 21), (0, 31)], dtype)


def test_fwd_many_items_per_cta_mixed_lengths():
    """The persistent CTAs walk through many work items each (heavy-first order, atomic hand-out): 2400 items of 1 to
    6 key-tile pairs on 148 CTAs, GQA; every (batch, head) is checked in full against the oracle on a second,
    smaller problem, and the large one through the causal prefix property against itself."""
    import starflashattention_b200 as sfa
    dtype = torch.bfloat16
    q = _rand((3, 1300, 8, 128), dtype, 41)
    k = _rand((3, 1300, 2, 128), dtype, 42)
    v = _rand((3, 1300, 2, 128), dtype, 43)
    o_ref, lse_ref, mag = _oracle(q, k, v, causal=True)
    o, lse = sfa.fwd(q.cuda(), k.cuda(), v.cuda(), causal=True)
    _check(o, lse, o_ref, lse_ref, dtype, "many items", mag=mag)
    # twice in a row on the same stream: the schedule counters must come back to zero by themselves
    o2, lse2 = sfa.fwd(q.cuda(), k.cuda(), v.cuda(), causal=True)
    assert torch.equal(o2, o) and torch.equal(lse2, lse)
    g = torch.Generator(device="cuda").manual_seed(44)
    qb, kb, vb = (torch.randn((50, 1536, 8, 128), generator=g, device="cuda", dtype=torch.float32).to(dtype) for _ in range(3))
    ob, lb = sfa.fwd(qb, kb, vb, causal=True)
    n = 640
    op, lp = sfa.fwd(qb[:, :n], kb[:, :n], vb[:, :n], causal=True)
    assert float((ob[:, :n].float() - op.float()).abs().max()) <= 2e-3 + 2.0 ** -8 * 4.0
    assert float((lb[:, :, :n] - lp).abs().max()) <= 1e-4


def test_fwd_graph_capture_and_two_streams():
    """The prefill call is capturable (its schedule counters belong to the captured launch) and two streams may run it
    at the same time (every launch has counters of its own)."""
    import starflashattention_b200 as sfa
    dtype = torch.float16
    q, k, v = (_rand((2, 512, 4, 128), dtype, s).cuda() for s in (51, 52, 53))
    o_ref, lse_ref = sfa.fwd(q, k, v, causal=True)
    torch.cuda.synchronize()
    out = torch.empty_like(q)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        sfa.fwd(q, k, v, causal=True, out=out, return_lse=False)
    for _ in range(3):
        out.zero_()
        graph.replay()
        torch.cuda.synchronize()
        assert torch.equal(out, o_ref)
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    outs = []
    for s in (s1, s2) * 4:
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            outs.append(sfa.fwd(q, k, v, causal=True)[0])
    torch.cuda.synchronize()
    assert all(torch.equal(x, o_ref) for x in outs)


def test_fwd_bf16_noise_floor():
    """the FlashAttention convention - the kernel's error against an fp64 evaluation must not exceed
    twice the error of a NAIVE evaluation in the data type (bf16 matmuls and softmax by torch on the GPU, the
    same formula as examples/python/testFlashDecoder.py:84-92) against the same fp64 result.  Printed with -s."""
    import starflashattention_b200 as sfa
    for dtype in (torch.bfloat16, torch.float16):
        q, k, v = (_rand((1, 1024, 4, 128), dtype, s) for s in (61, 62, 63))
        o64, _ = naive.attention_fwd(q, k, v, causal=True, acc_dtype=torch.float64)
        o, _ = sfa.fwd(q.cuda(), k.cuda(), v.cuda(), causal=True)
        qd, kd, vd = (t.cuda().transpose(1, 2) for t in (q, k, v))           # (B, H, S, D)
        s = torch.matmul(qd, kd.transpose(-1, -2)) / (128 ** 0.5)
        s = s.masked_fill(torch.ones(1024, 1024, device="cuda", dtype=torch.bool).triu(1), float("-inf"))
        o_naive = torch.matmul(torch.softmax(s, dim=-1), vd).transpose(1, 2)
        err_kernel = float((o.double().cpu() - o64).abs().max())
        err_naive = float((o_naive.double().cpu() - o64).abs().max())
        print(f"[noise floor] {dtype}: kernel max |O - O_fp64| = {err_kernel:.3e}, naive {dtype} = {err_naive:.3e}")
        assert err_kernel <= 2.0 * err_naive + 1e-5, (dtype, err_kernel, err_naive)


def test_fwd_full_c2_shape_spot_check():
    """BASELINE config 2 at full size (non-causal fp16 B=4 H=16 N=4096 d=64)."""
    import starflashattention_b200 as sfa
    dtype = torch.float16
    B, S, H, D = 4, 4096, 16, 64
    g = torch.Generator(device="cuda").manual_seed(2)
    q, k, v = (torch.randn((B, S, H, D), generator=g, device="cuda", dtype=torch.float32).to(dtype) for _ in range(3))
    o, lse = sfa.fwd(q, k, v)
    torch.cuda.synchronize()
    assert torch.isfinite(o).all() and torch.isfinite(lse).all()
    _spot_check(q, k, v, o, lse, False, [0, 1000, 4095], [(0, 0), (3, 15)], dtype)


@pytest.mark.parametrize("B,H,Hkv,chunks", [(4, 4, 2, None), (1, 8, 4, 3), (3, 2, 2, 2), (3, 8, 4, 12), (2, 4, 4, 5)])
def test_fwd_host_pipeline_matches_device_path(B, H, Hkv, chunks):
    """Pinned host tensors in, pinned host tensors out: the chunked three-stream pipeline must give exactly
    what one device-resident call gives (the units are independent), and match the oracle."""
    import starflashattention_b200 as sfa
    dtype = torch.float16
    q = _rand((B, 200, H, 128), dtype, 31).pin_memory()
    k = _rand((B, 333, Hkv, 128), dtype, 32).pin_memory()
    v = _rand((B, 333, Hkv, 128), dtype, 33).pin_memory()
    o_h, lse_h = sfa.fwd_host(q, k, v, causal=True, chunks=chunks)
    o_h2, lse_h2 = sfa.fwd_host(q, k, v, causal=True, chunks=chunks)      # second call reuses the staging buffers
    o_d, lse_d = sfa.fwd(q.cuda(), k.cuda(), v.cuda(), causal=True)
    torch.cuda.synchronize()
    assert o_h.is_pinned() and lse_h.is_pinned()
    assert torch.equal(o_h, o_d.cpu()) and torch.equal(lse_h, lse_d.cpu())
    assert torch.equal(o_h2, o_h) and torch.equal(lse_h2, lse_h)
    o_ref, lse_ref = naive.attention_fwd(q, k, v, causal=True)
    _check(o_h, lse_h, o_ref, lse_ref, dtype, "fwd_host")
    with pytest.raises(ValueError, match="pinned"):
        sfa.fwd_host(q.clone(), k, v)


@pytest.mark.parametrize("causal,Sq,Sk,chunks", [(True, 1000, 1000, 8), (True, 700, 1133, 6), (False, 600, 450, 6), (True, 520, 300, 9)])
def test_fwd_host_row_pieces_match_device_path(causal, Sq, Sk, chunks):
    """sfa_fwd_host cuts a batch entry into blocks of query rows; under the causal mask a block runs against the key
    prefix it can see (K/V travel incrementally).  Every row must come out exactly as in one device-resident call:
    same key tiles, same masks, same order."""
    import starflashattention_b200 as sfa
    dtype = torch.bfloat16
    q = _rand((2, Sq, 4, 128), dtype, 71).pin_memory()
    k = _rand((2, Sk, 2, 128), dtype, 72).pin_memory()
    v = _rand((2, Sk, 2, 128), dtype, 73).pin_memory()
    o_h, lse_h = sfa.fwd_host(q, k, v, causal=causal, chunks=chunks)
    o_d, lse_d = sfa.fwd(q.cuda(), k.cuda(), v.cuda(), causal=causal)
    torch.cuda.synchronize()
    assert torch.equal(o_h, o_d.cpu())
    assert torch.equal(lse_h, lse_d.cpu())
    o_ref, lse_ref, mag = _oracle(q, k, v, causal=causal)
    _check(o_h, lse_h, o_ref, lse_ref, dtype, "fwd_host rows", mag=mag)


@pytest.mark.parametrize("B,H,Hkv,S,causal", [(1, 16, 16, 2048, True), (5, 4, 2, 1536, True), (1, 32, 16, 1024, False), (2, 16, 16, 1300, True)])
@pytest.mark.parametrize("heavy", [0, 1])
def test_fwd_host_default_plan_matches_device_path(tuning, B, H, Hkv, S, causal, heavy):
    """The library's own plan (no explicit piece count): light calls get half head groups at the ends and geometrically
    smaller row blocks at the end of the call, heavy calls uniform head groups cut into row blocks of one wave of work
    items each, their kernels alternating between two streams; staging rings four pieces deep.  The byte thresholds are lowered so that these small shapes are cut the way C3 / C5 are; every row
    must still equal the single device-resident call bit for bit."""
    import starflashattention_b200 as sfa
    tuning("host_piece_bytes", 16 << 10, 8 << 20)
    tuning("host_taper_bytes", 32 << 10, 4 << 20)
    tuning("host_heavy", heavy, -1)      # 0: light kernels (C3-like plan), 1: heavy kernels (C5-like plan)
    tuning("host_wave_items", 8, 128)     # heavy plan: pieces of 8 work items instead of one wave of 128
    dtype = torch.bfloat16
    plan = sfa.fwd_host_plan(B, S, S, H, Hkv, 128, causal=causal)
    assert len(plan) >= 8
    if not heavy:
        assert len({p[4] - p[3] for p in plan}) >= 2                          # tapered: pieces of several sizes
    q = _rand((B, S, H, 128), dtype, 91).pin_memory()
    k = _rand((B, S, Hkv, 128), dtype, 92).pin_memory()
    v = _rand((B, S, Hkv, 128), dtype, 93).pin_memory()
    o_h, lse_h = sfa.fwd_host(q, k, v, causal=causal)
    assert sfa.fwd_host_last_pieces() == len(plan)
    o_h2, lse_h2 = sfa.fwd_host(q, k, v, causal=causal)                       # again: the rings are reused
    o_d, lse_d = sfa.fwd(q.cuda(), k.cuda(), v.cuda(), causal=causal)
    torch.cuda.synchronize()
    assert torch.equal(o_h, o_d.cpu()) and torch.equal(lse_h, lse_d.cpu())
    assert torch.equal(o_h2, o_h) and torch.equal(lse_h2, lse_h)
    assert sfa.fwd_watchdog() == 0


def test_fwd_host_head_slice_views():
    """A rank of the multi-GPU driver owns a range of heads of the host tensors: strided views in, strided view out
    (pitched DMAs); the other heads of the output tensor stay untouched."""
    import starflashattention_b200 as sfa
    dtype = torch.float16
    q = _rand((2, 300, 8, 128), dtype, 81).pin_memory()
    k = _rand((2, 300, 4, 128), dtype, 82).pin_memory()
    v = _rand((2, 300, 4, 128), dtype, 83).pin_memory()
    out = torch.full(q.shape, 7.0, dtype=dtype).pin_memory()
    o_h, lse_h = sfa.fwd_host(q[:, :, 2:6], k[:, :, 1:3], v[:, :, 1:3], causal=True, out=out[:, :, 2:6])
    torch.cuda.synchronize()
    o_ref, lse_ref = naive.attention_fwd(q[:, :, 2:6], k[:, :, 1:3], v[:, :, 1:3], causal=True)
    _check(out[:, :, 2:6], lse_h, o_ref, lse_ref, dtype, "head slice")
    assert bool((out[:, :, :2] == 7.0).all()) and bool((out[:, :, 6:] == 7.0).all())
