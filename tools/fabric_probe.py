#!/usr/bin/env python3
"""Run the fabric probes (NVLink P2P all-pairs, NCCL int32 allreduce) through the C ABI and print JSON.
Usage: python tools/fabric_probe.py [--ndev 8] [--bytes 1073741824] [--ar-count 268435456] [--out file]"""
import argparse
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ndev", type=int, default=0)
    ap.add_argument("--bytes", type=int, default=1 << 30)
    ap.add_argument("--ar-count", type=int, default=1 << 28)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--out", default="")
    ap.add_argument("--ar-sizes", default="", help="extra allreduce counts (log2 list) for a latency curve at max ndev")
    ap.add_argument("--modes", default="serial,serial_push,rounds,rounds_push,serial_ce,rounds_ce")
    ap.add_argument("--no-allreduce", action="store_true")
    args = ap.parse_args()
    gp = ge.load_package().gpuprobe
    nd = args.ndev or gp.device_count()
    res = {"ndev": nd, "bytes_per_pair": args.bytes}
    for name, mode in (("serial", gp.P2P_SERIAL), ("serial_push", gp.P2P_SERIAL_PUSH), ("rounds", gp.P2P_ROUNDS),
                       ("rounds_push", gp.P2P_ROUNDS_PUSH), ("serial_ce", gp.P2P_SERIAL_CE), ("rounds_ce", gp.P2P_ROUNDS_CE)):
        if name not in args.modes.split(","):
            continue
        rc, cells = gp.p2p_matrix(nd, args.bytes, iters=args.iters, warmup=3, mode=mode, raise_on_fail=False)
        g = [cells[s][t].gbps for s in range(nd) for t in range(nd) if s != t]
        res[f"p2p_{name}"] = {
            "rc": rc, "pairs": len(g), "min_gbps": round(min(g), 1), "median_gbps": round(statistics.median(g), 1),
            "max_gbps": round(max(g), 1), "frac_of_900_min": round(min(g) / 900, 4),
            "frac_of_900_median": round(statistics.median(g) / 900, 4),
            "all_verified": all(cells[s][t].verified for s in range(nd) for t in range(nd) if s != t),
            "matrix_gbps": [[round(cells[s][t].gbps, 1) for t in range(nd)] for s in range(nd)],
        }
    res["allreduce"] = []
    for k in (2, 4, 8):
        if k > nd or args.no_allreduce:
            continue
        ar = gp.allreduce_i32(k, args.ar_count, iters=20, warmup=5)
        res["allreduce"].append({"ndev": k, "bytes": ar.bytes, "mismatches": ar.mismatches, "ms_median": round(ar.ms_median, 4),
                                 "algbw_gbps": round(ar.algbw_gbps, 1), "busbw_gbps": round(ar.busbw_gbps, 1),
                                 "frac_of_900": round(ar.frac_of_peak, 4), "nccl_version": ar.nccl_version})
    res["allreduce_nvls_kernel"] = []
    for k in (2, 4, 8):
        if k > nd or args.no_allreduce:
            continue
        try:
            ar = gp.allreduce_i32_nvls(k, args.ar_count, iters=20, warmup=5)
            res["allreduce_nvls_kernel"].append({"ndev": k, "bytes": ar.bytes, "mismatches": ar.mismatches, "ms_median": round(ar.ms_median, 4),
                                                 "algbw_gbps": round(ar.algbw_gbps, 1), "busbw_gbps": round(ar.busbw_gbps, 1),
                                                 "frac_of_900": round(ar.frac_of_peak, 4)})
        except gp.GpuProbeError as e:
            res["allreduce_nvls_kernel"].append({"ndev": k, "error": str(e)})
    if args.ar_sizes:
        res["allreduce_curve"] = []
        for lg in [int(x) for x in args.ar_sizes.split(",")]:
            ar = gp.allreduce_i32(nd, 1 << lg, iters=20, warmup=5)
            res["allreduce_curve"].append({"ndev": nd, "bytes": ar.bytes, "us_median": round(ar.ms_median * 1e3, 2),
                                           "busbw_gbps": round(ar.busbw_gbps, 1), "mismatches": ar.mismatches})
    text = json.dumps(res)
    print(text)
    if args.out:
        with open(args.out, "w") as f:
            f.write(text + "\n")


if __name__ == "__main__":
    main()
