#!/usr/bin/env python3
"""Sweep vectorAdd kernel variants x sizes on one GPU through the C ABI; prints one JSON line per point.
Usage: python tools/vadd_sweep.py [--sizes 20,24,28,30] [--variants 1,2,...] [--iters 20] [--out file]"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", default="20,22,24,26,28,30")
    ap.add_argument("--variants", default="")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--dev", type=int, default=0)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    gp = ge.load_package().gpuprobe
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    measured = float(peaks.get("hbm_gbs", 0.0))
    variants = [int(v) for v in args.variants.split(",") if v] or list(range(1, gp.VADD_NUM_VARIANTS))
    out = open(args.out, "w") if args.out else None
    for lg in [int(s) for s in args.sizes.split(",")]:
        n = 1 << lg
        for v in variants:
            r = gp.vectoradd(n, dev=args.dev, iters=args.iters, warmup=args.warmup, variant=v, verify=1,
                             peak_measured_gbps=measured)
            line = json.dumps({"log2n": lg, "variant": gp.VARIANT_NAMES[v], "ms_best": round(r.ms_best, 5),
                               "ms_median": round(r.ms_median, 5), "gbps_best": round(r.gbps_best, 1),
                               "gbps_median": round(r.gbps_median, 1), "frac_nominal": round(r.frac_nominal, 4),
                               "frac_measured": round(r.frac_measured, 4), "l2_resident": r.l2_resident,
                               "sum64": r.sum64})
            print(line, flush=True)
            if out:
                out.write(line + "\n")
                out.flush()


if __name__ == "__main__":
    main()
