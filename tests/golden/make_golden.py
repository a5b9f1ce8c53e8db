#!/usr/bin/env python3
"""Generate tests/golden/golden_v1.json from the numpy restatement (oracle/oracle_np.py).

The reference holds no golden vectors for this path (parity
unpinned), so these fixtures are created here.  Run from the repo root:

    python tests/golden/make_golden.py

It refuses to write if the independent C oracle (oracle/gpuprobe_oracle.c)
disagrees with the numpy restatement on any fixture.
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as oc  # noqa: E402
from oracle import oracle_np as on  # noqa: E402

SEEDS = [0, 1, 2]
SIZES = [1, 3, 31, 50000, 1 << 20, (1 << 20) + 1]
KINDS = {"unit": 0, "edge": 1, "bits": 2}


def bits(x):
    return [int(v) for v in np.asarray(x).view(np.uint32)]


def main():
    g = {"version": 1, "spec": "DESIGN.md Generator/Digests", "hash": [], "vectoradd": [],
         "edge_table": None, "p2p": [], "allreduce": [], "offset_slices": []}

    for seed, stream, i in [(0, 0, 0), (0, 1, 0), (1, 0, 1), (2, 1, 49999), (0, 0, (1 << 32) - 1),
                            (0, 0, 1 << 32), (7, 0x103, 12345678901), (7, 0x207, (1 << 40) + 3)]:
        h = int(on.hash32(seed, stream, np.array([i], dtype=np.uint64))[0])
        assert h == oc.hash32(seed, stream, i)
        g["hash"].append({"seed": seed, "stream": stream, "i": i, "h": h})

    for kname, kind in KINDS.items():
        for seed in SEEDS:
            for n in SIZES:
                a, b = on.fill_inputs(n, seed, kind)
                c = on.vectoradd(a, b)
                assert not np.isnan(c).any()
                s, x = on.digest(c)
                assert (s, x) == oc.vectoradd_digest(n, seed, kind)
                ca, cb = oc.fill_inputs(n, seed, kind)
                assert oc.count_mismatch(oc.vectoradd(ca, cb), c) == 0
                g["vectoradd"].append({"kind": kname, "seed": seed, "n": n, "sum64": s, "xor32": x,
                                       "a_head": bits(a[:4]), "b_head": bits(b[:4]), "c_head": bits(c[:4]),
                                       "c_tail": bits(c[-2:])})

    # slices far from zero: exercises the 64-bit index path of the generator
    for first in [(1 << 32) - 3, (1 << 32), (3 << 31) + 11]:
        a, b = on.fill_inputs(64, 0, 0, first)
        c = on.vectoradd(a, b)
        assert on.digest(c) == oc.vectoradd_digest(64, 0, 0, first)
        g["offset_slices"].append({"first": first, "n": 64, "seed": 0, "kind": "unit",
                                   "sum64": on.digest(c)[0], "xor32": on.digest(c)[1], "c_head": bits(c[:4])})

    ea = np.repeat(on.EDGE_A, 16).view(np.float32)
    eb = np.tile(on.EDGE_B, 16).view(np.float32)
    ec = on.vectoradd(ea, eb)
    assert oc.count_mismatch(oc.vectoradd(np.ascontiguousarray(ea), np.ascontiguousarray(eb)), ec) == 0
    assert not np.isnan(ec).any()
    g["edge_table"] = {"a": bits(on.EDGE_A), "b": bits(on.EDGE_B), "c_row_major_a_by_b": bits(ec)}

    for seed in SEEDS:
        for src in (0, 3, 7):
            for nwords in (1, 1000, 1 << 18, (1 << 20) + 3):
                s, x = on.p2p_checksum(nwords, seed, src)
                assert (s, x) == oc.p2p_checksum(nwords, seed, src)
                g["p2p"].append({"seed": seed, "src": src, "nwords": nwords, "sum64": s, "xor32": x,
                                 "head": [int(v) for v in on.p2p_fill(min(nwords, 4), seed, src)]})

    for nranks in (2, 4, 8):
        for seed in SEEDS:
            e = on.allreduce_expected(4096, seed, nranks)
            assert (e == oc.allreduce_expected(4096, seed, nranks)).all()
            g["allreduce"].append({"seed": seed, "nranks": nranks, "n": 4096,
                                   "head": [int(v) for v in e[:8]],
                                   "sum64_of_bits": int(e.view(np.uint32).astype(np.uint64).sum())})

    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden_v1.json")
    with open(out, "w") as f:
        json.dump(g, f, indent=1)
        f.write("\n")
    print("wrote", out, os.path.getsize(out), "bytes")


if __name__ == "__main__":
    main()
