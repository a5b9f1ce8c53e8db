#!/usr/bin/env python3
"""Allreduce implementations side by side through the C ABI, one JSON line per (ranks, size).
Usage: python tools/nvls_probe.py [--counts 10,20,24,28] [--ndevs 2,4,8] [--impls nvls,two_shot,hybrid,push,auto,nccl]
Every result is compared with the CPU oracle's digest of the reduced buffer (`oracle_ok`)."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402
from oracle import oracle as oc  # noqa: E402  (a measurement tool, not the product: it may use the checker)

ap = argparse.ArgumentParser()
ap.add_argument("--counts", default="10,20,24,28")
ap.add_argument("--ndevs", default="")
ap.add_argument("--impls", default="nvls,two_shot,hybrid,push,copy_engine,nvls_ce,auto,nccl")
ap.add_argument("--iters", type=int, default=20)
ap.add_argument("--warmup", type=int, default=5)
ap.add_argument("--tag", default="")
args = ap.parse_args()
gp = ge.load_package().gpuprobe
n = gp.device_count()
ndevs = [int(x) for x in args.ndevs.split(",") if x] or [k for k in (2, 4, 8) if k <= n]
for nd in ndevs:
    for lg in [int(x) for x in args.counts.split(",")]:
        row = {"ndev": nd, "bytes": 4 << lg}
        if args.tag:
            row["tag"] = args.tag
        want = {}
        for impl in args.impls.split(","):
            try:
                r = gp.ALLREDUCE_IMPLS[impl](nd, 1 << lg, iters=args.iters, warmup=args.warmup)
                if r.rounds not in want:
                    want[r.rounds] = oc.allreduce_digest(1 << lg, 0, nd, rounds=r.rounds)
                row[impl] = {"us": round(r.ms_median * 1e3, 1), "busbw": round(r.busbw_gbps, 1), "algbw": round(r.algbw_gbps, 1),
                             "frac_of_900": round(r.frac_of_peak, 4), "mism": r.mismatches, "impl": gp.AR_IMPL_NAMES[r.impl],
                             "oracle_ok": (r.sum64, r.xor32) == want[r.rounds] and r.ranks_equal == 1}
            except gp.GpuProbeError as e:
                row[impl] = {"error": str(e)}
        print(json.dumps(row), flush=True)
