#!/usr/bin/env python3
"""bench.py -- the validator CUDA-workload path on B200, measured.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

One "step" = one pass of the hot path (C = A + B, fp32) over one batch of synthetic input:
n = 2^30 elements per array (BASELINE.json configs[1], the largest power-of-two sweep point
that leaves room for the e2e buffers; 12 GiB of algorithmic traffic per launch, far larger
than the 126 MB L2, so no flush is needed between iterations).

The JSON line (rank 0) follows the driver contract:
  value     whole-job GB/s (12 B/element) with A, B, C resident in HBM, CUDA-event timed,
            max over ranks; at N>1 every rank runs the same per-GPU workload (weak scaling,
            no data-path collective: vectorAdd does not exchange anything).
  e2e       same metric through the C ABI call a validator makes with HOST buffers
            (gpuprobe_vectoradd_buffers: H2D of A and B and D2H of C inside the timed region).
  roofline  the add kernel against the measured HBM copy peak (MEASURED_PEAKS.json).
  cpu_baseline  the oracle's host loop (a C port of the reference's CPU arithmetic) timed on
            this box's cores (N=1 only).
`--impl reference` times that host loop alone (rank 0), as the reference arm.
Only the cpu_baseline / reference legs touch oracle/; the measured product path is the CUDA
library behind the C ABI and fails loudly if it is missing.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

LOG2_N = 30
N_ELEMS = 1 << LOG2_N
BYTES_PER_ELEM = 12  # 2 x 4 B read + 4 B write
METRIC = "vectorAdd HBM GB/s"
UNIT = "GB/s"
# dram__bytes_read.sum + dram__bytes_write.sum per launch of the default kernel at n = 2^30,
# from the ncu --set full capture summarised in profiles/r01_vadd_ncu_full.md
NCU_TRAFFIC_BYTES_2P30 = 12_850_400_000
CPU_SAMPLE_LOG2 = 28


def workload_config(n_gpus: int) -> dict:
    return {
        "workload": f"vectorAdd fp32 C=A+B, n=2^{LOG2_N} elements per array per GPU "
                    f"(BASELINE.json configs[1] sweep point), bit-exact vs host loop",
        "n_elems_per_gpu": N_ELEMS,
        "algorithmic_bytes_per_elem": BYTES_PER_ELEM,
        "l2": "inputs (12 GiB per launch) are larger than L2; no flush needed",
        "parallelism": f"{n_gpus} independent per-GPU replicas, no collective on the data path",
    }


def load_peaks() -> tuple[float, str]:
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, burst copy)"
    except (OSError, KeyError, ValueError):
        return 6650.0, "fallback (B200_PROFILING.md: 6.65 TB/s)"


# ----------------------------------------------------------------------------- multi-rank arithmetic


def max_over_ranks(x: float, device: str = "cpu") -> float:
    """The slowest rank's time: every multi-GPU number is the max over ranks (no-op without a process group)."""
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()):
        return float(x)
    t = torch.tensor([x], device=device, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def whole_job_gbps(n_per_rank: int, world: int, steps: int, seconds: float) -> float:
    """Weak scaling: every rank processed n_per_rank elements per step; value = all ranks' bytes / slowest time."""
    return BYTES_PER_ELEM * n_per_rank * world * steps / seconds / 1e9


# ----------------------------------------------------------------------------- clocks


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 100 ms while the timed region runs."""

    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.samples: list[tuple[float, list[str]]] = []
        self.proc = None
        self.thread = None
        self.gpu_index = gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu_index), f"--query-gpu={self.FIELDS}",
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        self.thread = threading.Thread(target=self._pump, daemon=True)
        self.thread.start()

    def _pump(self):
        for line in self.proc.stdout:
            self.samples.append((time.time(), [x.strip() for x in line.split(",")]))

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()

    def summary(self, t0: float, t1: float, note: str) -> dict:
        rows = [s for (t, s) in self.samples if t0 <= t <= t1 and len(s) >= 7]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "window": note}
        sm = [float(r[0]) for r in rows if r[0].replace(".", "", 1).isdigit()]
        mx = [float(r[1]) for r in rows if r[1].replace(".", "", 1).isdigit()]
        pw = [float(r[2]) for r in rows if r[2].replace(".", "", 1).isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower() == "active" for r in rows)]
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "reasons": reasons, "samples": len(rows), "window": note}


def physical_gpu_index(local_rank: int) -> int:
    vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
    if vis:
        parts = [p for p in vis.split(",") if p.strip()]
        if local_rank < len(parts) and parts[local_rank].strip().isdigit():
            return int(parts[local_rank])
    return local_rank


# ----------------------------------------------------------------------------- reference arm


def cpu_host_loop_gbps(steps: int, warmup: int, sample_log2: int, min_seconds: float = 0.0):
    """Times the oracle's threaded host loop c[i] = a[i] + b[i] over a 2^sample_log2 sample."""
    import numpy as np

    from oracle import oracle as oc

    n = 1 << sample_log2
    threads = oc.host_threads()
    a, b = oc.fill_inputs(n, 0, oc.GEN_UNIT, threads=threads)  # the workload's first 2^sample_log2 elements
    c = np.empty_like(a)
    for _ in range(max(1, warmup)):
        oc.vectoradd_into(a, b, c, threads)
    times = []
    t_start = time.perf_counter()
    k = 0
    while k < steps or (time.perf_counter() - t_start) < min_seconds:
        t0 = time.perf_counter()
        oc.vectoradd_into(a, b, c, threads)
        times.append(time.perf_counter() - t0)
        k += 1
    total = sum(times)
    gbps = BYTES_PER_ELEM * n * len(times) / total / 1e9
    check = oc.digest(c[: 1 << 20]) == oc.vectoradd_digest(1 << 20, 0, oc.GEN_UNIT)
    # the same loop on ONE thread (SURVEY.md section 8d asks for both), three passes
    t1 = []
    for _ in range(3):
        t0 = time.perf_counter()
        oc.vectoradd_into(a, b, c, 1)
        t1.append(time.perf_counter() - t0)
    return {"gbps": gbps, "ms_per_step": 1e3 * total / len(times), "steps": len(times), "threads": threads,
            "n": n, "verified": bool(check), "gbps_1thread": BYTES_PER_ELEM * n / min(t1) / 1e9}


def run_reference(args, rank: int) -> int:
    if rank != 0:
        return 0
    r = cpu_host_loop_gbps(args.steps, args.warmup, CPU_SAMPLE_LOG2)
    sample = (f"each step = host loop over the first 2^{CPU_SAMPLE_LOG2} of the workload's 2^{LOG2_N} elements "
              f"({BYTES_PER_ELEM * r['n'] / 1e9:.2f} GB of traffic), {r['threads']} pthreads")
    line = {
        "impl": "reference", "metric": METRIC, "value": round(r["gbps"], 2), "unit": UNIT, "n_gpus": args.gpus,
        "steps": r["steps"], "warmup": args.warmup, "ms_per_step": round(r["ms_per_step"], 4),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.gpus),
        "cpu_baseline": {"value": round(r["gbps"], 2), "unit": UNIT, "cores": r["threads"], "kind": "port",
                         "sample": sample, "verified": r["verified"], "value_1thread": round(r["gbps_1thread"], 2)},
        "e2e": {"value": round(r["gbps"], 2), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "the reference's arithmetic for this path is a prebuilt third-party binary + Go host code, neither "
                "buildable here; this is the oracle's C port of the host loop on all host cores",
    }
    print(json.dumps(line), flush=True)
    return 0


# ----------------------------------------------------------------------------- our arm


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--e2e-steps", type=int, default=0, help="steps of the host-buffer leg (default min(steps, 10))")
    ap.add_argument("--no-extras", action="store_true", help="skip sweep / fabric extras")
    ap.add_argument("--variant", type=int, default=0)
    ap.add_argument("--skip-e2e", action="store_true", help="experiments only: leave out the host-buffer leg")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        return run_reference(args, rank)

    if args.warmup < 3:
        args.warmup = 3  # timing rule: at least 3 warm-up steps

    import numpy as np
    import torch
    import torch.distributed as dist

    import __graft_entry__ as ge

    gp = ge.load_package().gpuprobe  # raises if libgpuprobe.so is missing: no fallback
    assert torch.cuda.is_available(), "bench.py needs a GPU (there is no CPU fallback)"
    torch.cuda.set_device(local_rank)
    dev = local_rank
    distributed = world > 1
    gloo = None
    if distributed:
        dist.init_process_group(backend="nccl", device_id=torch.device("cuda", local_rank))
        gloo = dist.new_group(backend="gloo")

    def barrier():
        if distributed:
            dist.barrier()
        torch.cuda.synchronize()

    n = N_ELEMS
    stream = torch.cuda.current_stream().cuda_stream
    d_a = torch.empty(n, device="cuda", dtype=torch.float32)
    d_b = torch.empty(n, device="cuda", dtype=torch.float32)
    d_c = torch.empty(n, device="cuda", dtype=torch.float32)
    gp.fill_inputs_device(d_a, d_b, n, dev=dev, seed=0, gen_kind=gp.GEN_UNIT, stream=stream)
    d_c.fill_(float("nan"))
    torch.cuda.synchronize()

    # ---- device-resident leg: W warm-up, then exactly K timed steps --------------------
    sampler = ClockSampler(physical_gpu_index(local_rank))
    if rank == 0:
        sampler.start()
    for _ in range(args.warmup):
        gp.vectoradd_device(d_a, d_b, d_c, n, dev=dev, variant=args.variant, stream=stream)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t0_wall = time.time()
    ev0.record()
    for _ in range(args.steps):
        gp.vectoradd_device(d_a, d_b, d_c, n, dev=dev, variant=args.variant, stream=stream)
    ev1.record()
    barrier()
    t1_wall = time.time()
    ms_total = ev0.elapsed_time(ev1)
    launches = args.steps  # one kernel per step at this size (n is a multiple of every tile)

    clocks_note = "timed region"
    if rank == 0 and (t1_wall - t0_wall) < 0.5:
        # region too short for 100 ms sampling: run the same kernel untimed for ~1 s while sampling
        clocks_note = "extra ~1 s loop of the same kernel right after the timed region (region < 0.5 s)"
        t0_wall = time.time()
        while time.time() - t0_wall < 1.0:
            for _ in range(50):
                gp.vectoradd_device(d_a, d_b, d_c, n, dev=dev, variant=args.variant, stream=stream)
            torch.cuda.synchronize()
        t1_wall = time.time()
    clocks = None
    if rank == 0:
        time.sleep(0.15)
        sampler.stop()
        clocks = sampler.summary(t0_wall, t1_wall, clocks_note)

    ms_max = max_over_ranks(ms_total, "cuda")
    value = whole_job_gbps(n, world, args.steps, ms_max / 1e3)

    # parity of what was just timed: digest of C against the oracle-checked golden value
    sum64, xor32 = gp.digest_device(d_c, n, dev=dev, stream=stream)
    mism = gp.verify_device(d_c, n, dev=dev, seed=0, gen_kind=gp.GEN_UNIT, stream=stream)
    assert mism == 0, f"rank {rank}: {mism} elements differ from a+b"

    # ---- e2e leg: HOST buffers through the C ABI, copies inside the timed region ----------
    e2e_steps = args.e2e_steps or min(args.steps, 10)
    if args.skip_e2e:
        if rank == 0:
            used = gp.vectoradd(1 << 16, dev=dev, iters=1, warmup=0, variant=args.variant, verify=0).variant
            print(json.dumps({"experiment": True, "variant": gp.VARIANT_NAMES[used], "value": round(value, 1),
                              "ms_per_step": round(ms_max / args.steps, 5), "steps": args.steps, "clocks": clocks}), flush=True)
        if distributed:
            dist.destroy_process_group()
        return 0
    h_a, own_a = gp.pinned_array(n)
    h_b, own_b = gp.pinned_array(n)
    h_c, own_c = gp.pinned_array(n)
    torch.from_numpy(h_a).copy_(d_a)  # set-up, untimed: the same synthetic inputs, now in pinned host memory
    torch.from_numpy(h_b).copy_(d_b)
    h_c.fill(0)
    torch.cuda.synchronize()
    for _ in range(2):
        gp.vectoradd_buffers(h_a, h_b, h_c, n, dev=dev)
    barrier()
    te0 = time.perf_counter()
    for _ in range(e2e_steps):
        gp.vectoradd_buffers(h_a, h_b, h_c, n, dev=dev)  # returns after C is back in host memory
    torch.cuda.synchronize()
    te = time.perf_counter() - te0
    e2e_s = max_over_ranks(te, "cuda")
    e2e_value = whole_job_gbps(n, world, e2e_steps, e2e_s)
    # the host-side result is the same C, bit for bit
    hc_u = h_c.view(np.uint32)
    h_sum = int(hc_u.sum(dtype=np.uint64))
    assert h_sum == sum64, f"rank {rank}: e2e result digest {h_sum:#x} != device digest {sum64:#x}"
    del h_a, h_b, h_c, hc_u
    own_a.free(), own_b.free(), own_c.free()

    peak, peak_src = load_peaks()
    used_variant = gp.vectoradd(1 << 16, dev=dev, iters=1, warmup=0, variant=args.variant, verify=0).variant
    avg_launch_ms = ms_total / launches
    achieved = BYTES_PER_ELEM * n / (avg_launch_ms * 1e6)

    line = None
    if rank == 0:
        line = {
            "metric": METRIC, "value": round(value, 1), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms_max / args.steps, 5), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(world),
            "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": UNIT,
                         "frac": round(achieved / peak, 4), "traffic": NCU_TRAFFIC_BYTES_2P30,
                         "peak_source": peak_src, "kernel": f"gp::vadd_tile_kernel ({gp.VARIANT_NAMES[used_variant]})",
                         "algorithmic_bytes_per_launch": BYTES_PER_ELEM * n,
                         "avg_launch_ms": round(avg_launch_ms, 5),
                         "frac_of_nominal_8000": round(achieved / 8000.0, 4)},
            "e2e": {"value": round(e2e_value, 2), "unit": UNIT, "h2d_bytes_per_step": 8 * n, "d2h_bytes_per_step": 4 * n,
                    "steps": e2e_steps, "api": "gpuprobe_vectoradd_buffers (pinned host buffers)",
                    "ms_per_step": round(1e3 * e2e_s / e2e_steps, 3)},
            "gpu_launches": launches,
            "clocks": clocks,
            "parity": {"sum64": sum64, "xor32": xor32, "device_mismatches": mism, "e2e_host_digest_equal": True},
        }

    # ---- CPU baseline (N=1, rank 0): the oracle's host loop on this box's cores ---------
    if rank == 0 and world == 1:
        r = cpu_host_loop_gbps(steps=5, warmup=1, sample_log2=CPU_SAMPLE_LOG2, min_seconds=10.0)
        line["cpu_baseline"] = {
            "value": round(r["gbps"], 2), "unit": UNIT, "cores": r["threads"], "kind": "port",
            "sample": f"{r['steps']} passes of the host loop over the first 2^{CPU_SAMPLE_LOG2} of the workload's "
                      f"2^{LOG2_N} elements (~{r['steps'] * r['ms_per_step'] / 1e3:.0f} s of CPU work)",
            "verified": r["verified"], "value_1thread": round(r["gbps_1thread"], 2)}
        from oracle import oracle as oc

        line["parity"]["oracle_digest_first_2p28_equal"] = (
            gp.digest_device(d_c, 1 << 28, dev=dev, stream=stream) == oc.vectoradd_digest(1 << 28, 0, oc.GEN_UNIT))

    # ---- extras (informational; never part of `value`) ----------------------------------
    if not args.no_extras:
        del d_a, d_b, d_c
        torch.cuda.empty_cache()
        if rank == 0 and world == 1:
            sweep = []
            for lg in (14, 20, 24, 26, 28, 30, 32):
                nn = 50_000 if lg == 14 else 1 << lg
                r = gp.vectoradd(nn, dev=dev, iters=20, warmup=3, verify=1, peak_measured_gbps=peak)
                sweep.append({"n": nn, "us_median": round(r.ms_median * 1e3, 2), "us_back_to_back": round(r.ms_mean * 1e3, 2),
                              "gbps_median": round(r.gbps_median, 1),
                              "frac_measured": round(r.frac_measured, 4), "frac_nominal_8000": round(r.frac_nominal, 4),
                              "l2_resident": bool(r.l2_resident)})
            line["sweep"] = sweep
        if distributed:
            dist.barrier(group=gloo)
            if rank == 0:
                fabric = {}
                try:
                    rc, cells = gp.p2p_matrix(world, 1 << 30, iters=5, warmup=2, mode=gp.P2P_SERIAL, raise_on_fail=False)
                    g = [cells[s][t].gbps for s in range(world) for t in range(world) if s != t]
                    fabric["p2p_serial"] = {"rc": rc, "pairs": len(g), "bytes_per_pair": 1 << 30, "min_gbps": round(min(g), 1),
                                            "median_gbps": round(statistics.median(g), 1), "max_gbps": round(max(g), 1),
                                            "frac_of_900_min": round(min(g) / 900.0, 4),
                                            "all_verified": all(cells[s][t].verified for s in range(world) for t in range(world) if s != t)}
                    rc, cells = gp.p2p_matrix(world, 1 << 30, iters=5, warmup=2, mode=gp.P2P_ROUNDS, raise_on_fail=False)
                    g = [cells[s][t].gbps for s in range(world) for t in range(world) if s != t]
                    fabric["p2p_rounds"] = {"rc": rc, "min_gbps": round(min(g), 1), "median_gbps": round(statistics.median(g), 1),
                                            "max_gbps": round(max(g), 1)}
                    ar = gp.allreduce_i32(world, 1 << 28, iters=10, warmup=3)
                    fabric["allreduce_i32_nccl"] = {"ndev": world, "bytes": ar.bytes, "mismatches": ar.mismatches,
                                                    "busbw_gbps": round(ar.busbw_gbps, 1), "algbw_gbps": round(ar.algbw_gbps, 1),
                                                    "frac_of_900": round(ar.frac_of_peak, 4), "nccl_version": ar.nccl_version}
                    nv = gp.allreduce_i32_nvls(world, 1 << 28, iters=10, warmup=3)
                    fabric["allreduce_i32_nvls_kernel"] = {"ndev": world, "bytes": nv.bytes, "mismatches": nv.mismatches,
                                                           "busbw_gbps": round(nv.busbw_gbps, 1), "algbw_gbps": round(nv.algbw_gbps, 1),
                                                           "frac_of_900": round(nv.frac_of_peak, 4)}
                except gp.GpuProbeError as e:
                    fabric["error"] = str(e)
                line["fabric"] = fabric
            dist.barrier(group=gloo)

    if rank == 0:
        print(json.dumps(line), flush=True)
    if distributed:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
