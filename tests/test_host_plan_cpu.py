"""Piece plan of sfa_fwd_host (csrc/host_pipeline.cu::plan_fwd_host) - pure host logic, checked without a GPU:
every (batch entry, kv-head, query row) is covered exactly once, the pieces of a unit are consecutive and ascend, and
the default plan has the shape DESIGN.md section 5 describes (few large pieces; a light call gets a tapered tail)."""
import numpy as np
import pytest

import starflashattention_b200 as sfa


def _check_cover(plan, B, Sq, Hkv):
    seen = np.zeros((B, Hkv, Sq), dtype=np.int32)
    last_unit, last_r1, units_seen = -1, 0, set()
    for b, hk0, hkn, r0, r1, unit in plan:
        assert 0 <= b < B and 0 <= hk0 and hkn >= 1 and hk0 + hkn <= Hkv and 0 <= r0 < r1 <= Sq
        seen[b, hk0:hk0 + hkn, r0:r1] += 1
        if unit != last_unit:
            assert unit not in units_seen and r0 == 0          # a unit's pieces are consecutive and start at row 0
            units_seen.add(unit)
        else:
            assert r0 == last_r1                               # ... and ascend without gaps (K/V travel incrementally)
        last_unit, last_r1 = unit, r1
    assert (seen == 1).all()


@pytest.mark.parametrize("shape", [
    (8, 8192, 8192, 32, 32, 128, True), (1, 65536, 65536, 32, 32, 128, True), (1, 65536, 65536, 16, 16, 128, True),
    (1, 65536, 65536, 4, 4, 128, True), (8, 8192, 8192, 4, 4, 128, True), (4, 4096, 4096, 16, 16, 64, False),
    (2, 700, 700, 4, 2, 128, True), (3, 1000, 1333, 8, 4, 128, True), (1, 1, 5, 2, 1, 64, False),
])
@pytest.mark.parametrize("chunks", [None, 1, 5, 12])
def test_plan_covers_every_row_once(shape, chunks):
    B, Sq, Sk, Hq, Hkv, D, causal = shape
    plan = sfa.fwd_host_plan(B, Sq, Sk, Hq, Hkv, D, causal=causal, chunks=chunks)
    assert len(plan) >= 1
    _check_cover(plan, B, Sq, Hkv)


def test_default_plan_of_c3_and_c5():
    c3 = sfa.fwd_host_plan(8, 8192, 8192, 32, 32, 128)
    assert len(c3) <= 12
    assert all(r1 - r0 == 8192 for b, _, _, r0, r1, _ in c3 if b < 7)      # whole entries in the middle
    assert c3[-1][4] - c3[-1][3] <= 1024                                   # the exposed last piece is small
    # a heavy call (kernels as long as the copies) keeps uniform 4-head groups, cut into row blocks of one wave of work
    # items (4 heads x 8192 rows = 128 items for the 148 persistent CTAs)
    c5 = sfa.fwd_host_plan(1, 65536, 65536, 32, 32, 128)
    assert len(c5) == 64 and all(p[2] == 4 and p[4] - p[3] == 8192 for p in c5)
    c5_half = sfa.fwd_host_plan(1, 65536, 65536, 16, 16, 128)
    assert len(c5_half) == 32 and all(p[2] == 4 and p[4] - p[3] == 8192 for p in c5_half)
    c5_eighth = sfa.fwd_host_plan(1, 65536, 65536, 4, 4, 128)
    assert len(c5_eighth) == 8 and all(p[2] == 4 and p[4] - p[3] == 8192 for p in c5_eighth)


def test_explicit_piece_count_is_uniform_and_switch_turns_taper_off():
    plan = sfa.fwd_host_plan(3, 1000, 1000, 8, 4, 128, chunks=12)
    assert len(plan) == 12 and {p[4] - p[3] for p in plan} == {256, 232}
    sfa.tuning_set("host_taper", 0)
    try:
        flat = sfa.fwd_host_plan(8, 8192, 8192, 32, 32, 128)
    finally:
        sfa.tuning_set("host_taper", 1)
    assert len(flat) == 8 and all(p[3] == 0 and p[4] == 8192 for p in flat)


def test_plan_of_degenerate_dimensions_is_empty():
    for shape in [(0, 8, 8, 4, 4, 128), (2, 0, 8, 4, 4, 128), (2, 8, 8, 4, 4, 0), (2, 8, 8, 0, 4, 128), (2, 8, 8, 4, 3, 128)]:
        assert sfa.fwd_host_plan(*shape) == []
