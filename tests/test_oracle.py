"""CPU: pin oracle/naive.py against fixtures produced by the reference's own Python code
(tools/make_golden.py imports /root/reference/examples/python/testFlashDecoder.py)."""
import glob
import math
import os

import numpy as np
import pytest
import torch

from oracle import naive

DT = {"float16": torch.float16, "bfloat16": torch.bfloat16, "float32": torch.float32}


def _load(path):
    z = np.load(path, allow_pickle=False)
    B, H, D, L, M, cur, layer = [int(x) for x in z["meta"]]
    dt = DT[str(z["dtype"])]
    if str(z["fill"]) == "ones":
        qkv = torch.ones(B, 3, H, D, dtype=dt)
        kc = torch.ones(B, 1, cur - 1, H, D, dtype=dt)
        vc = torch.ones(B, 1, cur - 1, H, D, dtype=dt)
    else:
        qkv = torch.from_numpy(z["qkv"]).to(dt)
        kc = torch.from_numpy(z["k_cache_layer"]).to(dt)[:, None]
        vc = torch.from_numpy(z["v_cache_layer"]).to(dt)[:, None]
    return (B, H, D, cur), dt, qkv, kc, vc, torch.from_numpy(z["o_ref_fp32"]), torch.from_numpy(z["o_ref_native"])


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "ref_decode_*.npz"))),
                         ids=lambda p: os.path.basename(p)[11:-4])
def test_decode_step_matches_reference_python(path):
    (B, H, D, cur), dt, qkv, kc, vc, o32, onat = _load(path)
    seq = torch.full((B,), cur - 1, dtype=torch.int32)
    # the reference runs RoPE in the tensor dtype promoted by fp32 cos/sin and casts back (:45-58)
    o, lse, k_new, v_new = naive.decode_step_reference(qkv.float() if dt != torch.float32 else qkv,
                                                       kc.float(), vc.float(), seq, 0)
    ref = o32[:, :, 0, :]
    assert o.shape == ref.shape
    assert torch.allclose(o, ref, atol=2e-6, rtol=1e-5), float((o - ref).abs().max())
    # low-precision run of the reference itself stays within the north-star tolerance of the oracle
    o_lp, *_ = naive.decode_step_reference(qkv, kc, vc, seq, 0)
    tol = 2e-3 if dt != torch.bfloat16 else 2e-2
    assert float((o_lp.float() - onat[:, :, 0, :]).abs().max()) <= tol


def test_all_ones_fixture_is_one():
    # examples/cpp/testFlashDecoder.cc:63-78 — V == 1 so any valid softmax gives exactly 1
    qkv = torch.ones(2, 3, 32, 128, dtype=torch.float16)
    kc = torch.ones(2, 4, 512, 32, 128, dtype=torch.float16)
    seq = torch.full((2,), 511, dtype=torch.int32)
    o, lse, *_ = naive.decode_step_reference(qkv, kc, kc, seq, 0)
    assert torch.allclose(o, torch.ones_like(o), atol=1e-6)


@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("sq,sk", [(64, 64), (33, 97), (128, 40)])
def test_attention_fwd_against_explicit_loops(causal, sq, sk):
    if causal and sq > sk:
        pytest.skip("rows with no visible key are covered separately")
    torch.manual_seed(0)
    B, Hq, Hkv, D = 1, 4, 2, 16
    q = torch.randn(B, sq, Hq, D)
    k = torch.randn(B, sk, Hkv, D)
    v = torch.randn(B, sk, Hkv, D)
    o, lse = naive.attention_fwd(q, k, v, causal=causal, acc_dtype=torch.float64, chunk=16)
    for h in range(Hq):
        for i in range(0, sq, 7):
            lim = sk if not causal else i + (sk - sq) + 1
            s = [(float(q[0, i, h].double() @ k[0, j, h // 2].double()) / math.sqrt(D)) for j in range(lim)]
            m = max(s)
            w = [math.exp(x - m) for x in s]
            L = sum(w)
            ref = sum(wj * v[0, j, h // 2].double() for j, wj in enumerate(w)) / L
            assert torch.allclose(o[0, i, h], ref, atol=1e-12)
            assert abs(float(lse[0, h, i]) - (m + math.log(L))) < 1e-12


def test_causal_rows_without_keys():
    q = torch.randn(1, 8, 1, 16)
    k = torch.randn(1, 3, 1, 16)
    o, lse = naive.attention_fwd(q, k, k, causal=True)
    assert torch.all(o[0, :5] == 0) and torch.all(torch.isinf(lse[0, 0, :5]))
    assert torch.isfinite(lse[0, 0, 5:]).all()


def test_combine_splits_equals_unsplit():
    torch.manual_seed(1)
    q = torch.randn(1, 3, 32, dtype=torch.float64)
    k = torch.randn(1, 100, 3, 32, dtype=torch.float64)
    v = torch.randn(1, 100, 3, 32, dtype=torch.float64)
    o, lse = naive.decode_attention(q, k, v, torch.tensor([100]), acc_dtype=torch.float64)
    bounds = [0, 17, 17, 64, 100]                      # includes an empty split
    ops, ms, ls = [], [], []
    for a, b in zip(bounds[:-1], bounds[1:]):
        s = torch.einsum("hd,nhd->hn", q[0], k[0, a:b]) / math.sqrt(32)
        if b == a:
            ms.append(torch.full((3,), float("-inf"), dtype=torch.float64))
            ls.append(torch.zeros(3, dtype=torch.float64))
            ops.append(torch.zeros(3, 32, dtype=torch.float64))
            continue
        m = s.max(-1).values
        p = torch.exp(s - m[:, None])
        ms.append(m)
        ls.append(p.sum(-1))
        ops.append(torch.einsum("hn,nhd->hd", p, v[0, a:b]))
    O, L = naive.combine_splits(torch.stack(ops, 1), torch.stack(ms, 1), torch.stack(ls, 1))
    assert torch.allclose(O, o[0], atol=1e-12) and torch.allclose(L, lse[0], atol=1e-12)


def test_decode_zero_length():
    q = torch.randn(2, 2, 16)
    k = torch.randn(2, 8, 2, 16)
    o, lse = naive.decode_attention(q, k, k, torch.tensor([0, 5]))
    assert torch.all(o[0] == 0) and torch.isinf(lse[0]).all() and torch.isfinite(lse[1]).all()


def test_gather_pages_is_the_dense_cache_the_table_names():
    pool = torch.arange(6 * 4 * 2 * 8, dtype=torch.float32).reshape(6, 4, 2, 8)
    table = torch.tensor([[5, 0, 2], [1, 1, 3]], dtype=torch.int32)
    dense = naive.gather_pages(pool, table)
    assert dense.shape == (2, 12, 2, 8)
    for b in range(2):
        for i in range(3):
            assert torch.equal(dense[b, 4 * i:4 * i + 4], pool[int(table[b, i])])
