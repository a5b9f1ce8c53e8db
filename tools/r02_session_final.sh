#!/bin/bash
# Round-2 closing regression on a 2-GPU box: the whole GPU suite (also alone-in-process for the no-alloc file), smoke(), bench
# at N=1 (both arms) and N=2, and one ncu --set full capture of the bulk P2P kernel.
set -u
O=gpurun_out
mkdir -p $O
timeout 1500 python -m pytest tests -x -q -m gpu > $O/r02_final_pytest_gpu_2gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/r02_final_pytest_gpu_2gpu.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > $O/r02_final_pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?" | tee -a $O/r02_final_pytest_noalloc_alone.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/r02_final_smoke.log 2>&1; echo "smoke rc=$?"
CUDA_VISIBLE_DEVICES=0 timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > $O/r02_final_bench_1gpu_ref.json 2>/dev/null; echo "ref rc=$?"
CUDA_VISIBLE_DEVICES=0 timeout 900 python bench.py --steps 20 --warmup 3 > $O/r02_final_bench_1gpu.json 2> $O/r02_final_bench_1gpu.err; echo "bench1 rc=$?"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus 2 --steps 20 --warmup 3 > $O/r02_final_bench_2gpu.json 2> $O/r02_final_bench_2gpu.err; echo "bench2 rc=$?"
CMD="python tools/fabric_probe.py --ndev 2 --bytes 268435456 --iters 1 --modes serial --no-allreduce"
timeout 300 $CMD > $O/r02_p2p_plain_for_ncu_full.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:p2p_bulk -s 1 -c 1 -f -o $O/r02_p2p_bulk $CMD > $O/r02_p2p_ncu_full.log 2>&1; echo "ncu rc=$?"
echo done
