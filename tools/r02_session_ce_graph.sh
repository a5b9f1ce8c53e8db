#!/bin/bash
# Copy-engine two-shot replayed from a multi-GPU CUDA graph vs issued call by call; chunk sweep; NVLS + copy engines. Any GPU count >= 2.
set -u
O=gpurun_out
N=${1:-2}
mkdir -p $O
export GPUPROBE_NCCL_LIB=$(python -c "import importlib.util,os;s=importlib.util.find_spec('nvidia.nccl');print(os.path.join(list(s.submodule_search_locations)[0],'lib','libnccl.so.2'))")
timeout 900 python -m pytest tests/test_fabric_gpu.py -x -q -m gpu -k "copy_engine" > $O/r02_pytest_ce_graph_${N}gpu.log 2>&1; echo "pytest ce rc=$?" | tee -a $O/r02_pytest_ce_graph_${N}gpu.log
for g in 0 1; do for k in 1 2 4 8; do GPUPROBE_CE_DEBUG=1 GPUPROBE_CE_GRAPH=$g GPUPROBE_CE_CHUNKS=$k timeout 300 python tools/nvls_probe.py --ndevs $N --counts 28 --impls copy_engine --tag graph${g}_chunks$k >> $O/r02_ar_ce_graph_${N}gpu.jsonl 2>> $O/r02_ar_ce_graph_${N}gpu.err; done; done
for p in 300 500 700 850; do GPUPROBE_NVLSCE_PERMILLE=$p timeout 300 python tools/nvls_probe.py --ndevs $N --counts 28 --impls nvls_ce --tag permille$p >> $O/r02_ar_ce_graph_${N}gpu.jsonl 2>> $O/r02_ar_ce_graph_${N}gpu.err; done
timeout 300 python tools/nvls_probe.py --ndevs $N --counts 20,24,26,30 --impls copy_engine,auto --tag curve >> $O/r02_ar_ce_graph_${N}gpu.jsonl 2>> $O/r02_ar_ce_graph_${N}gpu.err
echo done
