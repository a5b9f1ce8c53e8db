#!/usr/bin/env python3
"""Sweep the host-buffer path's chunk size and slots (GPUPROBE_BUF_CHUNK_LOG2 / GPUPROBE_BUF_SLOTS). n = 2^28."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

gp = ge.load_package().gpuprobe
n = 1 << 28
a, oa = gp.pinned_array(n)
b, ob = gp.pinned_array(n)
c, oc_ = gp.pinned_array(n)
a[:] = 1.0
b[:] = 2.0
for lg in (20, 21, 22, 23, 24, 25):
    for slots in (2, 3, 4):
        os.environ["GPUPROBE_BUF_CHUNK_LOG2"] = str(lg)
        os.environ["GPUPROBE_BUF_SLOTS"] = str(slots)
        gp.vectoradd_buffers(a, b, c, n)
        t0 = time.perf_counter()
        for _ in range(5):
            gp.vectoradd_buffers(a, b, c, n)
        dt = (time.perf_counter() - t0) / 5
        assert c[0] == 3.0 and c[-1] == 3.0
        print(json.dumps({"chunk_mib": 4 << (lg - 20), "slots": slots, "ms": round(dt * 1e3, 2), "gbps": round(12 * n / dt / 1e9, 2)}), flush=True)
