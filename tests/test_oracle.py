"""The oracle against (1) hand-derived IEEE-754 known answers, (2) the committed golden
fixtures, (3) its independent numpy restatement.  CPU only.

The reference pins nothing at this boundary except a process exit code
(validator/main.go:1049, tests/scripts/checks.sh:124-128), so parity is "unpinned";
these are the anchors the repo supplies instead.
"""
import numpy as np
import pytest

from oracle import oracle as oc
from oracle import oracle_np as on

KINDS = {"unit": 0, "edge": 1, "bits": 2}

# (a_bits, b_bits, expected c_bits), each derived by hand from IEEE-754 binary32 RNE rules.
IEEE_KAT = [
    (0x3F800000, 0x3F800000, 0x40000000),  # 1 + 1 = 2
    (0x00000000, 0x80000000, 0x00000000),  # +0 + -0 = +0 under round-to-nearest
    (0x80000000, 0x80000000, 0x80000000),  # -0 + -0 = -0
    (0x3F800000, 0xBF800000, 0x00000000),  # x + (-x) = +0
    (0x00000001, 0x00000001, 0x00000002),  # denormals add exactly (no FTZ/DAZ)
    (0x007FFFFF, 0x00000001, 0x00800000),  # max denormal + min denormal = min normal
    (0x00800000, 0x80000001, 0x007FFFFF),  # min normal - min denormal = max denormal
    (0x3F7FFFFF, 0x33800000, 0x3F800000),  # (1 - 2^-24) + 2^-24 = 1 exactly
    (0x3F800000, 0x33800000, 0x3F800000),  # 1 + half-ulp: tie -> even (stays 1)
    (0x3F800001, 0x33800000, 0x3F800002),  # (1+ulp) + half-ulp: tie -> even (rounds up)
    (0x3F800001, 0x33000000, 0x3F800001),  # quarter ulp rounds down
    (0x3F800001, 0xBF800001, 0x00000000),  # exact cancellation
    (0x3F800001, 0xBF800000, 0x34000000),  # (1+2^-23) - 1 = 2^-23
    (0x7F7FFFFF, 0x7F7FFFFF, 0x7F800000),  # overflow -> +inf
    (0xFF7FFFFF, 0xFF7FFFFF, 0xFF800000),  # overflow -> -inf
    (0x7F7FFFFF, 0x73000000, 0x7F800000),  # max finite + half its ulp: tie -> even -> inf
    (0x7F7FFFFF, 0x72800000, 0x7F7FFFFF),  # max finite + quarter ulp stays
    (0x7F800000, 0xFF7FFFFF, 0x7F800000),  # inf + finite = inf
    (0xFF800000, 0x3F800000, 0xFF800000),  # -inf + finite = -inf
    (0x7F7FFFFF, 0xFF7FFFFF, 0x00000000),  # max - max = +0
    (0x3DCCCCCD, 0x3E4CCCCD, 0x3E99999A),  # 0.1f + 0.2f = 0.3f rounded (0.30000001192...)
]


def _f(bits):
    return np.array(bits, dtype=np.uint32).view(np.float32)


def test_fp_environment_is_ieee():
    assert oc.fp_env_ok(), "host FTZ/DAZ or non-nearest rounding would break bit parity"


@pytest.mark.parametrize("impl", ["c", "numpy"])
def test_ieee_known_answers(impl):
    a = _f([k[0] for k in IEEE_KAT])
    b = _f([k[1] for k in IEEE_KAT])
    want = np.array([k[2] for k in IEEE_KAT], dtype=np.uint32)
    got = (oc.vectoradd(a, b) if impl == "c" else on.vectoradd(a, b)).view(np.uint32)
    bad = [(hex(IEEE_KAT[i][0]), hex(IEEE_KAT[i][1]), hex(int(got[i])), hex(int(want[i])))
           for i in range(len(IEEE_KAT)) if got[i] != want[i]]
    assert not bad, bad


def test_hash_golden(golden):
    for h in golden["hash"]:
        assert oc.hash32(h["seed"], h["stream"], h["i"]) == h["h"]
        assert int(on.hash32(h["seed"], h["stream"], np.array([h["i"]], dtype=np.uint64))[0]) == h["h"]


def test_vectoradd_golden_digests(golden):
    for g in golden["vectoradd"]:
        kind = KINDS[g["kind"]]
        a, b = oc.fill_inputs(g["n"], g["seed"], kind)
        c = oc.vectoradd(a, b)
        assert list(a[:4].view(np.uint32)) == g["a_head"]
        assert list(b[:4].view(np.uint32)) == g["b_head"]
        assert list(c[:4].view(np.uint32)) == g["c_head"]
        assert list(c[-2:].view(np.uint32)) == g["c_tail"]
        assert oc.digest(c) == (g["sum64"], g["xor32"])
        # streaming digest (no arrays) and the threaded host loop agree too
        assert oc.vectoradd_digest(g["n"], g["seed"], kind) == (g["sum64"], g["xor32"])
        assert oc.count_mismatch(oc.vectoradd(a, b, threads=4), c) == 0
        # the sample's own tolerance check passes on every fixture
        assert oc.sample_check(a, b, c) == 0


def test_offset_slices_golden(golden):
    for g in golden["offset_slices"]:
        a, b = oc.fill_inputs(g["n"], g["seed"], 0, first=g["first"])
        c = oc.vectoradd(a, b)
        assert oc.digest(c) == (g["sum64"], g["xor32"])
        assert list(c[:4].view(np.uint32)) == g["c_head"]


def test_edge_table_golden(golden):
    t = golden["edge_table"]
    ea = np.repeat(np.array(t["a"], dtype=np.uint32), 16).view(np.float32)
    eb = np.tile(np.array(t["b"], dtype=np.uint32), 16).view(np.float32)
    c = oc.vectoradd(np.ascontiguousarray(ea), np.ascontiguousarray(eb))
    assert list(c.view(np.uint32)) == t["c_row_major_a_by_b"]
    assert not np.isnan(c).any()


def test_edge_generator_covers_the_table():
    a, b = oc.fill_inputs(1 << 14, 0, 1)
    pairs = set(zip(a.view(np.uint32).tolist(), b.view(np.uint32).tolist()))
    assert len(pairs) == 256
    assert not np.isnan(oc.vectoradd(a, b)).any()


def test_bits_generator_is_finite_with_denormals():
    a, b = oc.fill_inputs(1 << 18, 1, 2)
    for v in (a, b):
        assert np.isfinite(v).all()
        exp = (v.view(np.uint32) >> 23) & 0xFF
        assert (exp == 0).any(), "denormals must appear"
    assert not np.isnan(oc.vectoradd(a, b)).any()


def test_unit_generator_range():
    a, b = oc.fill_inputs(1 << 16, 2, 0)
    for v in (a, b):
        assert v.min() >= 0.0 and v.max() < 1.0


@pytest.mark.parametrize("kind", [0, 1, 2])
def test_c_and_numpy_restatements_agree(kind):
    for first in (0, 12345, (1 << 32) - 17, (1 << 35) + 1):
        a, b = oc.fill_inputs(4099, 5, kind, first)
        a2, b2 = on.fill_inputs(4099, 5, kind, first)
        assert oc.count_mismatch(a, np.ascontiguousarray(a2)) == 0
        assert oc.count_mismatch(b, np.ascontiguousarray(b2)) == 0
        assert oc.count_mismatch(oc.vectoradd(a, b), on.vectoradd(a2, b2)) == 0


def test_digest_is_chunk_composable_and_order_independent():
    n = 100003
    s, x = oc.vectoradd_digest(n, 1, 0)
    cuts = [0, 1, 4097, 50000, n]
    ss, xx = 0, 0
    for lo, hi in zip(cuts[:-1], cuts[1:]):
        cs, cx = oc.vectoradd_digest(hi - lo, 1, 0, first=lo)
        ss = (ss + cs) & (2**64 - 1)
        xx ^= cx
    assert (ss, xx) == (s, x)
    a, b = oc.fill_inputs(n, 1, 0)
    c = oc.vectoradd(a, b)
    perm = np.random.default_rng(0).permutation(n)
    assert oc.digest(np.ascontiguousarray(c[perm])) == (s, x)


def test_empty_input():
    a, b = oc.fill_inputs(0)
    assert oc.vectoradd(a, b).size == 0
    assert oc.digest(np.empty(0, dtype=np.float32)) == (0, 0)
    assert oc.vectoradd_digest(0) == (0, 0)


def test_commutative_bitwise():
    a, b = oc.fill_inputs(1 << 16, 3, 2)
    assert oc.count_mismatch(oc.vectoradd(a, b), oc.vectoradd(b, a)) == 0


def test_p2p_golden(golden):
    for g in golden["p2p"]:
        assert oc.p2p_checksum(g["nwords"], g["seed"], g["src"]) == (g["sum64"], g["xor32"])
        assert oc.p2p_fill(min(g["nwords"], 4), g["seed"], g["src"]).tolist() == g["head"]
        assert on.p2p_checksum(g["nwords"], g["seed"], g["src"]) == (g["sum64"], g["xor32"])


def test_allreduce_golden(golden):
    for g in golden["allreduce"]:
        e = oc.allreduce_expected(g["n"], g["seed"], g["nranks"])
        assert e[:8].tolist() == g["head"]
        assert int(e.view(np.uint32).astype(np.uint64).sum()) == g["sum64_of_bits"]
        # closed form = wraparound sum of the per-rank fills
        acc = np.zeros(g["n"], dtype=np.uint32)
        for r in range(g["nranks"]):
            acc += oc.allreduce_fill(g["n"], g["seed"], r).view(np.uint32)
        assert (acc.view(np.int32) == e).all()


def test_allreduce_digest_is_the_digest_of_the_expected_array(golden):
    """oracle_allreduce_digest (threaded, what the GPU parity tests and bench.py compare with) against the array form
    multiplied out in numpy, for every golden case, several round counts and a size that splits across threads."""
    for g in golden["allreduce"]:
        e = oc.allreduce_expected(g["n"], g["seed"], g["nranks"]).view(np.uint32)
        assert oc.allreduce_digest(g["n"], g["seed"], g["nranks"], rounds=1)[0] == g["sum64_of_bits"]
        for rounds in (1, 2, 5, 13):
            m = e * np.uint32(pow(g["nranks"], rounds - 1, 1 << 32))
            want = (int(m.sum(dtype=np.uint64)), int(np.bitwise_xor.reduce(m)))
            assert oc.allreduce_digest(g["n"], g["seed"], g["nranks"], rounds=rounds) == want
    n = (1 << 18) + 77
    e = oc.allreduce_expected(n, 7, 8, first=1 << 33).view(np.uint32) * np.uint32(pow(8, 3, 1 << 32))
    want = (int(e.sum(dtype=np.uint64)), int(np.bitwise_xor.reduce(e)))
    assert oc.allreduce_digest(n, 7, 8, rounds=4, first=1 << 33, threads=1) == want
    assert oc.allreduce_digest(n, 7, 8, rounds=4, first=1 << 33, threads=7) == want
    assert oc.allreduce_digest(0, 7, 8) == (0, 0)
