"""Independent numpy restatement of the oracle (second opinion for the C oracle).

TEST INFRASTRUCTURE ONLY -- PARITY UNPINNED (see oracle/gpuprobe_oracle.h).
Written from the specification in DESIGN.md ("Generator", "Digests"), not from
gpuprobe_oracle.c, so that the two restatements can check one another and the
golden fixtures in tests/golden/ (made by tests/golden/make_golden.py from this
module).  numpy float32 addition is IEEE-754 binary32 round-to-nearest-even, the
same single rounding as the host loop `c[i] = a[i] + b[i]`.
"""
from __future__ import annotations

import numpy as np

GEN_UNIT, GEN_EDGE, GEN_BITS = 0, 1, 2
STREAM_A, STREAM_B = 0, 1

_M32 = np.uint64(0xFFFFFFFF)

EDGE_A = np.array(
    [0x00000000, 0x80000000, 0x00000001, 0x80000001, 0x007FFFFF, 0x807FFFFF, 0x00800000, 0x80800000,
     0x3F7FFFFF, 0x3F800001, 0x3F800000, 0xBF800000, 0x7F7FFFFF, 0xFF7FFFFF, 0x7F800000, 0xFF800000],
    dtype=np.uint32)
EDGE_B = np.array(
    [0x00000000, 0x80000000, 0x00000001, 0x80000001, 0x007FFFFF, 0x807FFFFF, 0x00800000, 0x80800000,
     0x33800000, 0x33000000, 0xB3800000, 0xBF7FFFFF, 0xBF800001, 0x7F7FFFFF, 0xFF7FFFFF, 0x73000000],
    dtype=np.uint32)


def _mix32(x: np.ndarray) -> np.ndarray:
    """lowbias32 finaliser on uint64 carriers masked to 32 bits."""
    x = x & _M32
    x ^= x >> np.uint64(16)
    x = (x * np.uint64(0x7FEB352D)) & _M32
    x ^= x >> np.uint64(15)
    x = (x * np.uint64(0x846CA68B)) & _M32
    x ^= x >> np.uint64(16)
    return x


def hash32(seed: int, stream: int, idx: np.ndarray) -> np.ndarray:
    idx = np.asarray(idx, dtype=np.uint64)
    h0 = _mix32(np.array([(seed ^ (((stream + 1) * 0x9E3779B9) & 0xFFFFFFFF)) & 0xFFFFFFFF], dtype=np.uint64))[0]
    h1 = _mix32(h0 ^ (idx >> np.uint64(32)))
    h = _mix32(h1 ^ (idx & _M32))
    return h.astype(np.uint32)


def gen_bits(seed: int, stream: int, idx: np.ndarray, kind: int) -> np.ndarray:
    h = hash32(seed, stream, idx)
    if kind == GEN_EDGE:
        return (EDGE_A if stream == STREAM_A else EDGE_B)[h & np.uint32(15)]
    if kind == GEN_BITS:
        nan_exp = ((h >> np.uint32(23)) & np.uint32(0xFF)) == np.uint32(0xFF)
        return np.where(nan_exp, h ^ np.uint32(0x00800000), h).astype(np.uint32)
    f = (h >> np.uint32(8)).astype(np.float32) * np.float32(2.0 ** -24)
    return f.view(np.uint32)


def fill_inputs(n: int, seed: int = 0, kind: int = GEN_UNIT, first: int = 0):
    idx = np.arange(first, first + n, dtype=np.uint64)
    a = gen_bits(seed, STREAM_A, idx, kind).view(np.float32)
    b = gen_bits(seed, STREAM_B, idx, kind).view(np.float32)
    return a, b


def vectoradd(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore", invalid="ignore"):
        return (a.astype(np.float32) + b.astype(np.float32)).astype(np.float32)


def digest(c: np.ndarray):
    u = np.ascontiguousarray(c).view(np.uint32)
    s = int(u.astype(np.uint64).sum(dtype=np.uint64)) & 0xFFFFFFFFFFFFFFFF
    x = int(np.bitwise_xor.reduce(u)) if u.size else 0
    return s, x


def vectoradd_digest(n: int, seed: int = 0, kind: int = GEN_UNIT, first: int = 0, chunk: int = 1 << 22):
    s, x = 0, 0
    off = 0
    while off < n:
        m = min(chunk, n - off)
        a, b = fill_inputs(m, seed, kind, first + off)
        cs, cx = digest(vectoradd(a, b))
        s = (s + cs) & 0xFFFFFFFFFFFFFFFF
        x ^= cx
        off += m
    return s, x


def p2p_fill(nwords: int, seed: int, src: int, first: int = 0) -> np.ndarray:
    return hash32(seed, 0x100 + src, np.arange(first, first + nwords, dtype=np.uint64))


def p2p_checksum(nwords: int, seed: int, src: int, chunk: int = 1 << 22):
    s, x, off = 0, 0, 0
    while off < nwords:
        m = min(chunk, nwords - off)
        w = p2p_fill(m, seed, src, off)
        s = (s + int(w.astype(np.uint64).sum(dtype=np.uint64))) & 0xFFFFFFFFFFFFFFFF
        x ^= int(np.bitwise_xor.reduce(w))
        off += m
    return s, x


def allreduce_fill(n: int, seed: int, rank: int, first: int = 0) -> np.ndarray:
    return hash32(seed, 0x200 + rank, np.arange(first, first + n, dtype=np.uint64)).view(np.int32)


def allreduce_expected(n: int, seed: int, nranks: int, first: int = 0) -> np.ndarray:
    acc = np.zeros(n, dtype=np.uint32)
    for r in range(nranks):
        acc = acc + allreduce_fill(n, seed, r, first).view(np.uint32)  # uint32 wraps mod 2^32
    return acc.view(np.int32)
