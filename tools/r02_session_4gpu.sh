#!/bin/bash
# Round-2 4-GPU session: copy-engine allreduce (chunk sweep), NVLS + copy engines (split sweep), P2P copy-engine modes,
# fabric tests for the new paths, then bench.py as the driver runs it at N=4.
set -u
O=gpurun_out
mkdir -p $O
export GPUPROBE_NCCL_LIB=$(python -c "import importlib.util,os;s=importlib.util.find_spec('nvidia.nccl');print(os.path.join(list(s.submodule_search_locations)[0],'lib','libnccl.so.2'))")
for k in 1 2 4; do GPUPROBE_CE_CHUNKS=$k timeout 300 python tools/nvls_probe.py --ndevs 4 --counts 28 --impls copy_engine --tag chunks$k >> $O/r02_ar_ce_4gpu.jsonl 2>> $O/r02_ar_ce_4gpu.err; done
for p in 300 500 650 800; do GPUPROBE_NVLSCE_PERMILLE=$p timeout 300 python tools/nvls_probe.py --ndevs 4 --counts 28 --impls nvls_ce --tag permille$p >> $O/r02_ar_ce_4gpu.jsonl 2>> $O/r02_ar_ce_4gpu.err; done
timeout 300 python tools/nvls_probe.py --ndevs 4,3 --counts 28 --impls hybrid,nvls,push,copy_engine,nccl --tag same_box >> $O/r02_ar_ce_4gpu.jsonl 2>> $O/r02_ar_ce_4gpu.err
timeout 300 python tools/fabric_probe.py --ndev 4 --no-allreduce --iters 5 --modes serial,rounds,rounds_push,serial_ce,rounds_ce --out $O/r02_p2p_4gpu.json > /dev/null 2> $O/r02_p2p_4gpu.err; echo "p2p rc=$?"
timeout 600 python -m pytest tests/test_fabric_gpu.py -x -q -m gpu -k "copy_engine or auto" > $O/r02_pytest_ce_4gpu.log 2>&1; echo "pytest ce rc=$?" | tee -a $O/r02_pytest_ce_4gpu.log
unset GPUPROBE_NCCL_LIB
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 4 --steps 20 --warmup 3 > $O/r02_bench_4gpu.json 2> $O/r02_bench_4gpu.err; echo "bench4 rc=$?"
echo done
