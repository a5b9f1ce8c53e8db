#!/bin/bash
# Round-2, fourth 2-GPU session: copy-engine two-shot allreduce and copy-engine P2P modes (parity, then bandwidth).
set -u
O=gpurun_out
mkdir -p $O
export GPUPROBE_NCCL_LIB=$(python -c "import importlib.util,os;s=importlib.util.find_spec('nvidia.nccl');print(os.path.join(list(s.submodule_search_locations)[0],'lib','libnccl.so.2'))")
timeout 900 python -m pytest tests/test_fabric_gpu.py -x -q -m gpu -k "copy_engine" > $O/r02_pytest_ce_2gpu.log 2>&1; echo "pytest ce rc=$?" | tee -a $O/r02_pytest_ce_2gpu.log
for k in 1 2 4 8 16; do GPUPROBE_CE_CHUNKS=$k timeout 300 python tools/nvls_probe.py --ndevs 2 --counts 28 --impls copy_engine --tag chunks$k >> $O/r02_ar_ce_2gpu.jsonl 2>> $O/r02_ar_ce_2gpu.err; done
timeout 300 python tools/nvls_probe.py --ndevs 2 --counts 10,20,24,26,28,30 --impls copy_engine,push,nccl --tag curve >> $O/r02_ar_ce_2gpu.jsonl 2>> $O/r02_ar_ce_2gpu.err
timeout 300 python tools/fabric_probe.py --ndev 2 --no-allreduce --iters 5 --modes serial_ce,rounds_ce,rounds_push --out $O/r02_p2p_ce_2gpu.json > /dev/null 2> $O/r02_p2p_ce_2gpu.err; echo "p2p ce rc=$?"
echo done
