"""The oracle's arithmetic against the IEEE-754 standard itself.

Parity for this path is unpinned in the reference (no golden vectors exist there), so the oracle's host loop
`c[i] = a[i] + b[i]` -- executed by the host FPU -- is checked here against oracle/ieee754_add.py, a binary32
adder written from the standard in pure integer arithmetic (unpack, align with guard/round/sticky, add or
subtract, normalise, round to nearest even, pack).  Millions of operand pairs from every regime: the probe's
three generators, near-cancellation, the subnormal range, the overflow boundary, and the 16 x 16 edge table.
CPU only.
"""
import numpy as np
import pytest

from oracle import ieee754_add as ie
from oracle import oracle as oc
from oracle import oracle_np as on


def _host_add(ua, ub):
    a = np.ascontiguousarray(ua.astype(np.uint32)).view(np.float32)
    b = np.ascontiguousarray(ub.astype(np.uint32)).view(np.float32)
    return oc.vectoradd(a, b).view(np.uint32)


def _finite(u):
    u = u.astype(np.uint32)
    return np.where(((u >> 23) & 0xFF) == 0xFF, u ^ np.uint32(0x00800000), u).astype(np.uint32)


def _assert_same(ua, ub):
    want, got = _host_add(ua, ub), ie.f32_add_bits(ua, ub)
    bad = np.nonzero(want != got)[0]
    assert bad.size == 0, [(hex(int(ua[i])), hex(int(ub[i])), hex(int(want[i])), hex(int(got[i]))) for i in bad[:5]]


def test_edge_table_matches_integer_adder(golden):
    ea = np.repeat(on.EDGE_A, 16)
    eb = np.tile(on.EDGE_B, 16)
    got = ie.f32_add_bits(ea, eb)
    assert got.tolist() == golden["edge_table"]["c_row_major_a_by_b"]
    _assert_same(ea, eb)


@pytest.mark.parametrize("kind", [0, 1, 2])
def test_generator_outputs_match_integer_adder(kind):
    for seed in (0, 1):
        a, b = oc.fill_inputs(1 << 20, seed, kind)
        _assert_same(a.view(np.uint32), b.view(np.uint32))


def test_near_cancellation():
    rng = np.random.default_rng(1)
    ua = _finite(rng.integers(0, 2**32, size=1 << 20, dtype=np.uint64))
    delta = rng.integers(-64, 65, size=ua.size)
    ub = _finite(((ua.astype(np.int64) + delta) & 0xFFFFFFFF).astype(np.uint32) ^ np.uint32(0x80000000))
    _assert_same(ua, ub)


def test_subnormal_range_both_signs():
    rng = np.random.default_rng(2)
    n = 1 << 20
    ua = rng.integers(0, 1 << 25, size=n, dtype=np.uint64).astype(np.uint32) | (rng.integers(0, 2, size=n, dtype=np.uint64).astype(np.uint32) << 31)
    ub = rng.integers(0, 1 << 25, size=n, dtype=np.uint64).astype(np.uint32) | (rng.integers(0, 2, size=n, dtype=np.uint64).astype(np.uint32) << 31)
    _assert_same(ua, ub)


def test_overflow_boundary():
    rng = np.random.default_rng(3)
    n = 1 << 20
    ua = _finite(rng.integers(0, 1 << 24, size=n, dtype=np.uint64).astype(np.uint32) | np.uint32(0x7F000000))
    ub = _finite(rng.integers(0, 1 << 24, size=n, dtype=np.uint64).astype(np.uint32) | np.uint32(0x7E800000))
    got = ie.f32_add_bits(ua, ub)
    assert (got == 0x7F800000).any() and (got != 0x7F800000).any()  # the sample straddles the boundary
    _assert_same(ua, ub)
    _assert_same(ua | np.uint32(0x80000000), ub | np.uint32(0x80000000))


def test_wide_exponent_gaps_and_ties():
    """Exponent differences 0..60 with low bits chosen to land exactly on rounding ties."""
    rows_a, rows_b = [], []
    for d in range(0, 61):
        for ea in (1, 2, 64, 127, 200, 254):
            eb = ea - d
            if eb < 0:
                continue
            for fa in (0, 1, 0x400000, 0x7FFFFF, 0x7FFFFE):
                for fb in (0, 1, 0x400000, 0x7FFFFF, 0x000001 << min(d, 22)):
                    for sgn in (0, 0x80000000):
                        rows_a.append((ea << 23) | fa)
                        rows_b.append(((eb << 23) | (fb & 0x7FFFFF)) | sgn)
    ua, ub = _finite(np.array(rows_a, dtype=np.uint64)), _finite(np.array(rows_b, dtype=np.uint64))
    _assert_same(ua, ub)


def test_domain_is_enforced():
    with pytest.raises(ValueError):
        ie.f32_add_bits(np.array([0x7FC00000], dtype=np.uint32), np.array([0], dtype=np.uint32))
    with pytest.raises(ValueError):
        ie.f32_add_bits(np.array([0x7F800000], dtype=np.uint32), np.array([0xFF800000], dtype=np.uint32))
