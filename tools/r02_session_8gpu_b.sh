#!/bin/bash
# Round-2, second 8-GPU session: copy-engine allreduce and P2P modes at 4 and 8 ranks, then bench.py as the driver runs it at N=8.
set -u
O=gpurun_out
mkdir -p $O
export GPUPROBE_NCCL_LIB=$(python -c "import importlib.util,os;s=importlib.util.find_spec('nvidia.nccl');print(os.path.join(list(s.submodule_search_locations)[0],'lib','libnccl.so.2'))")
for k in 1 2 3 4; do GPUPROBE_CE_CHUNKS=$k timeout 300 python tools/nvls_probe.py --ndevs 8,4 --counts 28 --impls copy_engine --tag chunks$k >> $O/r02_ar_ce_8gpu.jsonl 2>> $O/r02_ar_ce_8gpu.err; done
for p in 500 650 800 900; do GPUPROBE_NVLSCE_PERMILLE=$p timeout 300 python tools/nvls_probe.py --ndevs 8,4 --counts 28 --impls nvls_ce --tag permille$p >> $O/r02_ar_ce_8gpu.jsonl 2>> $O/r02_ar_ce_8gpu.err; done
GPUPROBE_NVLS_CTAS=2 GPUPROBE_NVLSCE_PERMILLE=800 timeout 300 python tools/nvls_probe.py --ndevs 8 --counts 28 --impls nvls_ce --tag permille800_ctas2 >> $O/r02_ar_ce_8gpu.jsonl 2>> $O/r02_ar_ce_8gpu.err
timeout 300 python tools/nvls_probe.py --ndevs 8 --counts 28 --impls hybrid,nvls --tag same_box >> $O/r02_ar_ce_8gpu.jsonl 2>> $O/r02_ar_ce_8gpu.err
timeout 300 python tools/nvls_probe.py --ndevs 3,5,6 --counts 28 --impls copy_engine,nvls,nccl --tag odd_rank_counts >> $O/r02_ar_ce_8gpu.jsonl 2>> $O/r02_ar_ce_8gpu.err
timeout 300 python tools/fabric_probe.py --ndev 8 --no-allreduce --iters 5 --modes serial_ce,rounds_ce --out $O/r02_p2p_ce_8gpu.json > /dev/null 2> $O/r02_p2p_ce_8gpu.err; echo "p2p ce rc=$?"
timeout 600 python -m pytest tests/test_fabric_gpu.py -x -q -m gpu -k "copy_engine or auto" > $O/r02_pytest_ce_8gpu.log 2>&1; echo "pytest ce rc=$?" | tee -a $O/r02_pytest_ce_8gpu.log
unset GPUPROBE_NCCL_LIB
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 bench.py --impl reference --gpus 8 --steps 5 --warmup 1 > $O/r02_bench_8gpu_ref.json 2> $O/r02_bench_8gpu_ref.err; echo "bench8 ref rc=$?"
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus 8 --steps 20 --warmup 3 > $O/r02_bench_8gpu.json 2> $O/r02_bench_8gpu.err; echo "bench8 rc=$?"
echo done
