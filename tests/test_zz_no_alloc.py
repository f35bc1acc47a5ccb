"""sfa_decode_static (include/star_flash_attn.h): the split-KV decode entry point that allocates nothing - its
per-split partials live in a static device array.  Called through the C ABI and compared with sfa_decode on the same
inputs, twice in a row and with another shape in between (the array keeps the previous call's contents).  Collected
last on purpose."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _rand(shape, dtype, seed):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(shape, generator=g).to(dtype)


def _bound(ref):   # the decode tests' tolerance (tests/test_decode_gpu.py)
    return 2e-3 + 4e-3 * ref.float().abs()


def _both(q, kc, vc, seqlens, num_splits):
    """(o, lse) of sfa_decode and of sfa_decode_static on the same device tensors."""
    from starflashattention_b200 import _capi, interface
    outs = []
    for fn in (_capi.lib.sfa_decode, _capi.lib.sfa_decode_static):
        o = torch.full_like(q, float("nan"))
        lse = torch.full(q.shape[:2], float("nan"), dtype=torch.float32, device=q.device)
        a = interface._decode_args(q, kc, vc, seqlens, o, lse, None, num_splits)
        free0 = torch.cuda.mem_get_info()[0]
        _capi.check(fn(a, interface._stream()))
        torch.cuda.synchronize()
        outs.append((o, lse, free0 - torch.cuda.mem_get_info()[0]))
    return outs


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_decode_static_matches_decode_on_consecutive_calls(dtype):
    from starflashattention_b200 import _capi
    assert _capi.lib.sfa_decode_static_scratch_bytes() == 16 << 20
    cases = [
        # B, Hq, Hkv, D, M, lens, splits: split counts above 8 take the partials + combine path (the scratch)
        (2, 8, 2, 128, 3000, [3000, 1111], 16),
        (2, 8, 2, 128, 3000, [3000, 1111], 16),      # the same call again
        (3, 4, 4, 64, 900, [900, 1, 0], 11),         # another shape over the previous call's leftovers, an empty sequence
        (1, 32, 8, 128, 8192, [8192], 0),            # the launcher's own split count (few units, long cache)
        (2, 8, 2, 128, 3000, [3000, 1111], 16),
    ]
    for i, (B, Hq, Hkv, D, M, lens, splits) in enumerate(cases):
        q = _rand((B, Hq, D), dtype, 10 + i).cuda()
        kc = _rand((B, M, Hkv, D), dtype, 20 + i).cuda()
        vc = _rand((B, M, Hkv, D), dtype, 30 + i).cuda()
        seqlens = torch.tensor(lens, dtype=torch.int32).cuda()
        (o_ref, lse_ref, _), (o, lse, grew) = _both(q, kc, vc, seqlens, splits)
        assert not torch.isnan(o).any(), f"case {i}: output rows left unwritten"
        err = (o.float() - o_ref.float()).abs()
        assert bool((err <= _bound(o_ref)).all()), f"case {i}: max err {float(err.max())}"
        assert torch.equal(o, o_ref), f"case {i}: same kernels, same splits - expected identical bits"
        fin = torch.isfinite(lse_ref)
        assert torch.equal(torch.isfinite(lse), fin)
        assert float((lse[fin] - lse_ref[fin]).abs().max()) <= 1e-3
        if i > 0:   # (the first call may load the module lazily)
            assert grew <= 0, f"case {i}: device memory in use grew by {grew} bytes across sfa_decode_static"


def test_decode_static_rejects_partials_beyond_the_array():
    from starflashattention_b200 import _capi, interface
    dtype = torch.float16
    B, Hq, Hkv, D, M = 64, 32, 8, 128, 512
    q = _rand((B, Hq, D), dtype, 1).cuda()
    kc = _rand((B, M, Hkv, D), dtype, 2).cuda()
    vc = _rand((B, M, Hkv, D), dtype, 3).cuda()
    o = torch.zeros_like(q)
    a = interface._decode_args(q, kc, vc, None, o, None, None, 32)      # 64 * 32 * 32 rows * 520 B = 34 MB
    rc = _capi.lib.sfa_decode_static(a, interface._stream())
    assert rc == _capi.SFA_ERR_INVALID
    msg = _capi.lib.sfa_last_error().decode()
    assert "static array holds 16777216" in msg and "need" in msg
    torch.cuda.synchronize()
    assert float(o.abs().max()) == 0.0                                   # nothing was launched
    a = interface._decode_args(q, kc, vc, None, o, None, None, 4)       # 4 splits fit (and merge in a cluster)
    _capi.check(_capi.lib.sfa_decode_static(a, interface._stream()))
    o_ref, _ = interface.decode(q, kc, vc, num_splits=4)
    torch.cuda.synchronize()
    assert torch.equal(o, o_ref)
