#!/usr/bin/env python3
"""NVLS (multimem) allreduce kernel vs NCCL through the C ABI. Usage: python tools/nvls_probe.py [--counts 20,24,28]"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--counts", default="10,20,24,28")
ap.add_argument("--ndevs", default="")
ap.add_argument("--skip-nccl", action="store_true")
args = ap.parse_args()
gp = ge.load_package().gpuprobe
n = gp.device_count()
ndevs = [int(x) for x in args.ndevs.split(",") if x] or [k for k in (2, 4, 8) if k <= n]
for nd in ndevs:
    for lg in [int(x) for x in args.counts.split(",")]:
        row = {"ndev": nd, "bytes": 4 << lg}
        try:
            a = gp.allreduce_i32_nvls(nd, 1 << lg, iters=20, warmup=5)
            row["nvls"] = {"us": round(a.ms_median * 1e3, 1), "busbw": round(a.busbw_gbps, 1), "algbw": round(a.algbw_gbps, 1), "mism": a.mismatches}
        except gp.GpuProbeError as e:
            row["nvls"] = {"error": str(e)}
        try:
            c = gp.allreduce_i32_p2p(nd, 1 << lg, iters=20, warmup=5)
            row["p2p2shot"] = {"us": round(c.ms_median * 1e3, 1), "busbw": round(c.busbw_gbps, 1), "mism": c.mismatches}
        except gp.GpuProbeError as e:
            row["p2p2shot"] = {"error": str(e)}
        try:
            h = gp.allreduce_i32_hybrid(nd, 1 << lg, iters=20, warmup=5)
            row["hybrid"] = {"us": round(h.ms_median * 1e3, 1), "busbw": round(h.busbw_gbps, 1), "mism": h.mismatches}
        except gp.GpuProbeError as e:
            row["hybrid"] = {"error": str(e)}
        if not args.skip_nccl:
            b = gp.allreduce_i32(nd, 1 << lg, iters=20, warmup=5)
            row["nccl"] = {"us": round(b.ms_median * 1e3, 1), "busbw": round(b.busbw_gbps, 1), "mism": b.mismatches}
        print(json.dumps(row), flush=True)
