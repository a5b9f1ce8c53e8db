#!/bin/bash
# Round-2, fifth 2-GPU session: the copy-engine engine after its refactor into ce_twoshot.cuh, and the NVLS + copy-engine hybrid.
set -u
O=gpurun_out
mkdir -p $O
export GPUPROBE_NCCL_LIB=$(python -c "import importlib.util,os;s=importlib.util.find_spec('nvidia.nccl');print(os.path.join(list(s.submodule_search_locations)[0],'lib','libnccl.so.2'))")
timeout 900 python -m pytest tests/test_fabric_gpu.py -x -q -m gpu -k "copy_engine or auto" > $O/r02_pytest_ce2_2gpu.log 2>&1; echo "pytest ce rc=$?" | tee -a $O/r02_pytest_ce2_2gpu.log
timeout 300 python tools/nvls_probe.py --ndevs 2 --counts 28 --impls copy_engine,nvls_ce,push,nvls --tag refactor >> $O/r02_ar_ce2_2gpu.jsonl 2>> $O/r02_ar_ce2_2gpu.err
for p in 200 400 600; do GPUPROBE_NVLSCE_PERMILLE=$p timeout 300 python tools/nvls_probe.py --ndevs 2 --counts 28 --impls nvls_ce --tag permille$p >> $O/r02_ar_ce2_2gpu.jsonl 2>> $O/r02_ar_ce2_2gpu.err; done
echo done
