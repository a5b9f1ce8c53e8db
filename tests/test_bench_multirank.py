"""The N>1 path of bench.py on CPU: two gloo ranks on 127.0.0.1 (no GPU).
vectorAdd does not shard and has no data-path collective, so the only multi-rank logic is
(1) the max-over-ranks timing, (2) the whole-job (weak-scaling) value, (3) the reference arm's
rank gating under torchrun (rank 0 alone works and prints, the others exit 0)."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import json, os, sys
sys.path.insert(0, os.environ["BENCH_ROOT"])
import torch.distributed as dist
import bench
dist.init_process_group(backend="gloo")
rank, world = dist.get_rank(), dist.get_world_size()
ms_local = 10.0 + 5.0 * rank           # rank 1 is the slow one
ms_max = bench.max_over_ranks(ms_local, "cpu")
value = bench.whole_job_gbps(1 << 20, world, 7, ms_max / 1e3)
every = bench.gather_over_ranks(ms_local, "cpu")   # per-rank times, in rank order, on every rank
dist.barrier()
print(json.dumps({"rank": rank, "world": world, "ms_max": ms_max, "value": value, "every": every}), flush=True)
dist.destroy_process_group()
'''


def _free_port():
    import socket

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _torchrun(args, env_extra=None, timeout=300):
    env = dict(os.environ, BENCH_ROOT=ROOT, OMP_NUM_THREADS="1")
    env.update(env_extra or {})
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), *args]
    return subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=timeout, cwd=ROOT)


def test_max_over_ranks_and_whole_job_value_gloo_world2(tmp_path):
    w = tmp_path / "worker.py"
    w.write_text(WORKER)
    r = _torchrun([str(w)])
    assert r.returncode == 0, r.stderr[-2000:]
    import re

    # two ranks share one stdout: their lines can land on the same line, so pick the JSON objects out by shape
    rows = [json.loads(m) for m in re.findall(r"\{[^{}]*\}", r.stdout)]
    assert all(x["every"] == [10.0, 15.0] for x in rows)  # what per_gpu_gbps at N > 1 is computed from
    assert sorted(x["rank"] for x in rows) == [0, 1]
    for x in rows:
        assert x["world"] == 2
        assert x["ms_max"] == 15.0                       # the slower rank's time, on both ranks
        want = 12 * (1 << 20) * 2 * 7 / 0.015 / 1e9      # both ranks' bytes over the slowest time
        assert abs(x["value"] - want) < 1e-9 * want


def test_single_process_helpers_need_no_process_group():
    sys.path.insert(0, ROOT)
    import bench

    assert bench.max_over_ranks(3.5) == 3.5
    assert bench.gather_over_ranks(3.5) == [3.5]
    assert bench.whole_job_gbps(1 << 30, 1, 1, 1.0) == pytest.approx(12 * (1 << 30) / 1e9)


def test_reference_arm_under_torchrun_prints_once_from_rank0():
    """Driver contract: `bench.py --impl reference` is launched like our arm; rank 0 alone runs and prints."""
    r = _torchrun(["bench.py", "--impl", "reference", "--gpus", "2", "--steps", "2", "--warmup", "1"])
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["n_gpus"] == 2 and d["unit"] == "GB/s" and d["value"] > 0
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["verified"] is True
    assert d["e2e"] == {"value": d["value"], "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
