#!/bin/bash
# Round-2 fabric session on a 2-GPU box: parity tests, NVLink counters, bidirectional P2P shapes, allreduce variants,
# NCCL algorithm record. Everything lands in gpurun_out/.
set -u
O=gpurun_out
mkdir -p $O
export GPUPROBE_NCCL_LIB=$(python -c "import importlib.util,os;s=importlib.util.find_spec('nvidia.nccl');print(os.path.join(list(s.submodule_search_locations)[0],'lib','libnccl.so.2'))")
nvidia-smi topo -m > $O/r02_topo_2gpu.txt 2>&1
timeout 900 python -m pytest tests/test_fabric_gpu.py -x -q -m gpu > $O/r02_pytest_fabric_2gpu.log 2>&1; echo "pytest fabric rc=$?" | tee -a $O/r02_pytest_fabric_2gpu.log
# link traffic vs algorithmic bytes: per-link data counters around one SERIAL matrix (2 pairs x (3 warm-up + 10) x 1 GiB)
nvidia-smi nvlink -gt d > $O/r02_nvlink_before.txt 2>&1
timeout 300 python tools/fabric_probe.py --ndev 2 --modes serial --no-allreduce --out $O/r02_nvlink_probe.json > /dev/null 2>$O/r02_nvlink_probe.err
nvidia-smi nvlink -gt d > $O/r02_nvlink_after.txt 2>&1
timeout 600 tools/exp/bin/p2p_bidir_exp 5 > $O/r02_p2p_bidir.jsonl 2> $O/r02_p2p_bidir.err; echo "bidir rc=$?"
timeout 600 python tools/nvls_probe.py --ndevs 2 --counts 20,28 > $O/r02_ar_2gpu.jsonl 2> $O/r02_ar_2gpu.err; echo "ar rc=$?"
GPUPROBE_NVLS_ST_V4=1 timeout 300 python tools/nvls_probe.py --ndevs 2 --counts 28 --impls nvls,hybrid --tag st_v4 >> $O/r02_ar_2gpu.jsonl 2>> $O/r02_ar_2gpu.err
# NCCL: which algorithm / protocol / channel count it picks here, then forced choices
for cfg in "default" "NCCL_ALGO=NVLS" "NCCL_ALGO=Ring" "NCCL_NVLS_ENABLE=0" "NCCL_ALGO=Ring NCCL_MIN_NCHANNELS=32"; do
  echo "=== $cfg" >> $O/r02_nccl_trials.log
  ( [ "$cfg" != default ] && export $cfg; NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT,ENV,TUNING timeout 300 python tools/nvls_probe.py --ndevs 2 --counts 28 --impls nccl --tag "$cfg" 2>&1 | grep -E "^\{|NVLS|Algo|algo|Channel|channels|nvls|NCCL_|Connected|busbw" | head -60 ) >> $O/r02_nccl_trials.log 2>&1
done
echo done
