#!/usr/bin/env python3
"""Headless full node-validation pass (BASELINE.json configs[4]): driver -> toolkit -> cuda -> plugin, each as
the `nvidia-validator` mirror with the reference's env, writing *-ready files into a fresh OUTPUT_DIR; reports
time-to-ready per component and in total.  The Kubernetes hops (pod scheduling, image pull, the 5 s phase poll)
are replaced by a child process and are therefore excluded from the time -- say so next to the number.
Usage: python tools/full_pass.py [--vadd-elems N] [--p2p-bytes B] [--ar-count C] [--poll-ms 50] [--json]"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
VDIR = os.path.join(ROOT, "gpu-operator_b200", "validator")


def full_pass(vadd_elems=0, p2p_bytes=0, ar_count=0, poll_ms=50, out_dir=None):
    out = out_dir or tempfile.mkdtemp(prefix="validations-")
    base = {"PATH": os.environ["PATH"], "OUTPUT_DIR": out, "NODE_NAME": os.uname().nodename, "OPERATOR_NAMESPACE": "gpu-operator",
            "GPUPROBE_MANIFEST_DIR": os.path.join(VDIR, "manifests"), "GPUPROBE_BIN_DIR": os.path.join(VDIR, "bin"),
            "GPUPROBE_POLL_MS": str(poll_ms), "VALIDATOR_IMAGE": "headless", "GPUPROBE_HEADLESS": "1", "LD_LIBRARY_PATH": os.environ.get("LD_LIBRARY_PATH", "")}
    base.update({k: v for k, v in os.environ.items() if k.startswith("GPUPROBE_") and k not in base})
    if vadd_elems:
        base["GPUPROBE_VADD_ELEMS"] = str(vadd_elems)
    if p2p_bytes:
        base["GPUPROBE_P2P_BYTES"] = str(p2p_bytes)
    if ar_count:
        base["GPUPROBE_AR_COUNT"] = str(ar_count)
    steps = []
    t_start = time.perf_counter()
    # env exactly as the DaemonSet's init containers set it (assets/state-operator-validation/0500_daemonset.yaml:29-142)
    for comp, extra in (("driver", {}), ("toolkit", {}), ("cuda", {}), ("plugin", {"WITH_WORKLOAD": "true"})):
        env = dict(base, COMPONENT=comp, **extra)
        t0 = time.perf_counter()
        r = subprocess.run([os.path.join(VDIR, "bin", "nvidia-validator")], env=env, capture_output=True, text=True)
        steps.append({"component": comp, "rc": r.returncode, "seconds": round(time.perf_counter() - t0, 3),
                      "stderr_tail": r.stderr.strip().splitlines()[-1:] if r.returncode else []})
        if r.returncode:
            break
    total = time.perf_counter() - t_start
    ready = sorted(f for f in os.listdir(out) if f.endswith("-ready"))
    results = None
    try:
        with open(os.path.join(out, "gpuprobe-results.json")) as f:
            results = json.load(f)
    except (OSError, ValueError):
        pass
    return {"output_dir": out, "steps": steps, "ready_files": ready, "time_to_ready_s": round(total, 3),
            "excluded": "pod scheduling, image pull, kubelet container start, 5 s pod-phase poll quantum",
            "results": results}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--vadd-elems", type=int, default=0)
    ap.add_argument("--p2p-bytes", type=int, default=0)
    ap.add_argument("--ar-count", type=int, default=0)
    ap.add_argument("--poll-ms", type=int, default=50)
    a = ap.parse_args()
    res = full_pass(a.vadd_elems, a.p2p_bytes, a.ar_count, a.poll_ms)
    print(json.dumps(res))
    sys.exit(0 if all(s["rc"] == 0 for s in res["steps"]) and len(res["steps"]) == 4 else 1)
