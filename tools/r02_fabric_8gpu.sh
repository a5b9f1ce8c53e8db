#!/bin/bash
# Round-2 fabric session on the 8-GPU box: allreduce variants at 4 and 8 ranks, P2P matrix in all modes (bulk and
# register-path kernels), NCCL algorithm record, then the whole fabric test file. Everything lands in gpurun_out/.
set -u
O=gpurun_out
mkdir -p $O
export GPUPROBE_NCCL_LIB=$(python -c "import importlib.util,os;s=importlib.util.find_spec('nvidia.nccl');print(os.path.join(list(s.submodule_search_locations)[0],'lib','libnccl.so.2'))")
timeout 600 python tools/nvls_probe.py --ndevs 8,4 --counts 28 > $O/r02_ar_8gpu.jsonl 2> $O/r02_ar_8gpu.err; echo "ar rc=$?"
timeout 300 python tools/nvls_probe.py --ndevs 8 --counts 10,20,24 --impls auto,nccl,push --tag curve >> $O/r02_ar_8gpu.jsonl 2>> $O/r02_ar_8gpu.err
GPUPROBE_NVLS_ST_V4=1 timeout 300 python tools/nvls_probe.py --ndevs 8,4 --counts 28 --impls nvls,hybrid --tag st_v4 >> $O/r02_ar_8gpu.jsonl 2>> $O/r02_ar_8gpu.err
timeout 600 python tools/fabric_probe.py --ndev 8 --no-allreduce --iters 5 --out $O/r02_fabric_8gpu.json > /dev/null 2> $O/r02_fabric_8gpu.err; echo "p2p rc=$?"
GPUPROBE_P2P_KERNEL=ldst timeout 600 python tools/fabric_probe.py --ndev 8 --no-allreduce --iters 5 --modes rounds,rounds_push --out $O/r02_fabric_8gpu_ldst.json > /dev/null 2>> $O/r02_fabric_8gpu.err
for cfg in "default" "NCCL_ALGO=NVLS" "NCCL_ALGO=NVLSTree" "NCCL_ALGO=Ring"; do
  echo "=== $cfg" >> $O/r02_nccl_trials_8gpu.log
  ( [ "$cfg" != default ] && export $cfg; NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT,ENV,TUNING timeout 300 python tools/nvls_probe.py --ndevs 8 --counts 28 --impls nccl --tag "$cfg" 2>&1 | grep -E "^\{|AllReduce:|NVLS comm|nvls channels|set by environment|NCCL version" | sort | uniq -c | sort -rn | head -12 ) >> $O/r02_nccl_trials_8gpu.log 2>&1
done
timeout 900 python -m pytest tests/test_fabric_gpu.py -x -q -m gpu > $O/r02_pytest_fabric_8gpu.log 2>&1; echo "pytest fabric rc=$?" | tee -a $O/r02_pytest_fabric_8gpu.log
echo done
