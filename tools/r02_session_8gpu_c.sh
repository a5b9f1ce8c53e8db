#!/bin/bash
# Round-2, third 8-GPU session: copy-engine two-shot with a capped number of concurrent copy streams, alone and beside the multimem kernel.
set -u
O=gpurun_out
mkdir -p $O
for ns in 1 2 3 4 7; do GPUPROBE_CE_STREAMS=$ns timeout 200 python tools/nvls_probe.py --ndevs 8 --counts 28 --impls copy_engine --tag streams$ns >> $O/r02_ar_ce_streams_8gpu.jsonl 2>> $O/r02_ar_ce_streams_8gpu.err; done
for ns in 1 2; do for p in 700 850 930; do GPUPROBE_CE_STREAMS=$ns GPUPROBE_NVLSCE_PERMILLE=$p timeout 200 python tools/nvls_probe.py --ndevs 8 --counts 28 --impls nvls_ce --tag streams${ns}_permille$p >> $O/r02_ar_ce_streams_8gpu.jsonl 2>> $O/r02_ar_ce_streams_8gpu.err; done; done
GPUPROBE_CE_STREAMS=2 timeout 200 python tools/nvls_probe.py --ndevs 5,6 --counts 28 --impls copy_engine,nvls --tag streams2 >> $O/r02_ar_ce_streams_8gpu.jsonl 2>> $O/r02_ar_ce_streams_8gpu.err
timeout 300 python -m pytest tests/test_fabric_gpu.py -x -q -m gpu -k "copy_engine" > $O/r02_pytest_ce_streams_8gpu.log 2>&1; echo "pytest rc=$?"
echo done
