#!/bin/bash
# Round-2, second 2-GPU session: the whole GPU test suite, bench.py at N=2 as the driver launches it, and NVLink byte
# counters (ncu nvlrx/nvltx, user vs protocol bytes) for one pull and one push launch of the P2P kernel.
set -u
O=gpurun_out
mkdir -p $O
export GPUPROBE_NCCL_LIB=$(python -c "import importlib.util,os;s=importlib.util.find_spec('nvidia.nccl');print(os.path.join(list(s.submodule_search_locations)[0],'lib','libnccl.so.2'))")
timeout 1200 python -m pytest tests -x -q -m gpu > $O/r02_pytest_gpu_2gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/r02_pytest_gpu_2gpu.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 20 --warmup 3 > $O/r02_bench_2gpu.json 2> $O/r02_bench_2gpu.err; echo "bench2 rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29518 bench.py --impl reference --gpus 2 --steps 5 --warmup 1 > $O/r02_bench_2gpu_ref.json 2>> $O/r02_bench_2gpu.err; echo "bench2 ref rc=$?"
M=nvlrx__bytes.sum,nvlrx__bytes_data_user.sum,nvlrx__bytes_data_protocol.sum,nvltx__bytes.sum,nvltx__bytes_data_user.sum,nvltx__bytes_data_protocol.sum,nvlrx__bytes_packet_request.sum,nvlrx__bytes_packet_response.sum,nvltx__bytes_packet_request.sum,nvltx__bytes_packet_response.sum,gpu__time_duration.sum
CMD="python tools/fabric_probe.py --ndev 2 --bytes 268435456 --iters 1 --modes serial,serial_push --no-allreduce"
timeout 300 $CMD > $O/r02_p2p_plain_for_ncu.log 2>&1 && \
timeout 600 ncu --metrics $M --clock-control none -k regex:p2p_bulk -c 16 --csv --log-file $O/r02_p2p_nvlink_ncu.csv $CMD > $O/r02_p2p_ncu.log 2>&1; echo "ncu rc=$?"
echo done
