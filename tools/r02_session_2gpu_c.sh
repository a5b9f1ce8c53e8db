#!/bin/bash
# Round-2, third 2-GPU session: the NCCL-static build of the library (what the validator image ships) through the
# NCCL allreduce tests, the copy engines as a reference point for the P2P figures, and the per-call cost of the
# host-buffer entry point at the reference's 50,000 elements with and without the cached staging.
set -u
O=gpurun_out
mkdir -p $O
GPUPROBE_LIB=$PWD/gpu-operator_b200/lib/libgpuprobe_ncclstatic.so GPUPROBE_NCCL_LIB=/nonexistent/libnccl.so.2 timeout 600 python -m pytest tests/test_fabric_gpu.py -x -q -m gpu -k "allreduce_int32_bit_exact or nvls_and_nccl or auto_allreduce" > $O/r02_pytest_ncclstatic_2gpu.log 2>&1; echo "ncclstatic rc=$?" | tee -a $O/r02_pytest_ncclstatic_2gpu.log
GPUPROBE_LIB=$PWD/gpu-operator_b200/lib/libgpuprobe_ncclstatic.so python -c "
import __graft_entry__ as g, json
gp=g.load_package().gpuprobe
r=gp.allreduce_i32(2, 1<<28, iters=10, warmup=3)
print(json.dumps({'lib': gp.LIB_PATH.split('/')[-1], 'nccl_version': r.nccl_version, 'busbw_gbps': round(r.busbw_gbps,1), 'mismatches': r.mismatches, 'ranks_equal': r.ranks_equal}))
" >> $O/r02_pytest_ncclstatic_2gpu.log 2>&1
P2P_EXP_CE_ONLY=1 timeout 300 tools/exp/bin/p2p_bidir_exp 5 > $O/r02_p2p_copy_engine.jsonl 2> $O/r02_p2p_copy_engine.err; echo "ce rc=$?"
for cache in 1 0; do
GPUPROBE_BUF_CACHE=$cache python - >> $O/r02_buffers_call_latency.jsonl 2>> $O/r02_buffers_call_latency.err <<'PY'
import json, os, time
import numpy as np
import __graft_entry__ as g
gp = g.load_package().gpuprobe
n = 50_000
a = np.random.default_rng(0).random(n, dtype=np.float32); b = a[::-1].copy(); c = np.empty_like(a)
def per_call_us(fn, reps=300):
    fn(); t0 = time.perf_counter()
    for _ in range(reps): fn()
    return round(1e6 * (time.perf_counter() - t0) / reps, 1)
row = {"n": n, "GPUPROBE_BUF_CACHE": os.environ["GPUPROBE_BUF_CACHE"], "pageable_us": per_call_us(lambda: gp.vectoradd_buffers(a, b, c, n))}
assert (c == a + b).all()
pa, oa = gp.pinned_array(n); pb, ob = gp.pinned_array(n); pc, oc_ = gp.pinned_array(n)
pa[:], pb[:] = a, b
row["pinned_us"] = per_call_us(lambda: gp.vectoradd_buffers(pa, pb, pc, n))
import torch
w = torch.zeros(1 << 20, dtype=torch.int32, device="cuda")
row["digest_1Mi_words_us"] = per_call_us(lambda: gp.digest_device(w, 1 << 20))
row["digest_noalloc_1Mi_words_us"] = per_call_us(lambda: gp.digest_device_noalloc(w, 1 << 20))
print(json.dumps(row))
PY
done
echo done
