#!/bin/bash
# Round-2 single-GPU session: GPU tests, smoke, bench (both arms), vectorAdd phase-separation / tile-order experiments,
# ncu launch list of the bench command and one ncu --set full capture of the hot kernel.
set -u
O=gpurun_out
mkdir -p $O
timeout 1200 python -m pytest tests -x -q -m gpu > $O/r02_pytest_gpu_1gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/r02_pytest_gpu_1gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/r02_smoke.log 2>&1; echo "smoke rc=$?"
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > $O/r02_bench_1gpu_ref.json 2> $O/r02_bench_1gpu_ref.err; echo "ref rc=$?"
timeout 900 python bench.py --steps 20 --warmup 3 > $O/r02_bench_1gpu.json 2> $O/r02_bench_1gpu.err; echo "bench rc=$?"
timeout 900 python bench.py > $O/r02_bench_1gpu_default.json 2> $O/r02_bench_1gpu_default.err; echo "bench default rc=$?"
for f in vadd_t256_u1 phased swz burst; do timeout 300 tools/exp/bin/vadd_exp 30 10 $f >> $O/r02_exp_vadd_phase.jsonl 2>> $O/r02_exp_vadd_phase.err; done; echo "exp rc=$?"
CMD="python bench.py --steps 5 --warmup 3 --no-full-pass"
timeout 600 $CMD > $O/r02_bench_plain_for_ncu.json 2> $O/r02_bench_plain_for_ncu.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r02_bench_launches.csv $CMD > $O/r02_bench_under_ncu.log 2>&1; echo "ncu list rc=$?"
CMD2="python bench.py --steps 5 --warmup 3 --no-extras --skip-e2e"
timeout 600 $CMD2 > $O/r02_bench_plain_for_ncu2.json 2> $O/r02_bench_plain_for_ncu2.err && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:vadd_tile -s 3 -c 3 -o $O/r02_vadd_u1 -f $CMD2 > $O/r02_ncu_full.log 2>&1; echo "ncu full rc=$?"
echo done
