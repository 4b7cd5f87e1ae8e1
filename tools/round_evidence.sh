#!/bin/bash
# round-end evidence on one B200: GPU parity tests, bench lines, ncu launch list, ncu full set of the three pipeline
# kernels (break and continue); then, here:  python tools/ncu_summaries.py <tag> cfg4 break
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/tests_gpu.log 2>&1; echo "tests rc=$?"
timeout 600 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
for c in cfg1 cfg2 cfg3 cfg5; do timeout 400 python bench.py --config $c --steps 100 --warmup 5 > gpurun_out/bench_$c.json 2> gpurun_out/bench_$c.err; echo "$c rc=$?"; done
S="python tools/ncu_step.py 64 break 4"
$S > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $S > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_insert|k_mid|k_gather" -s 6 -c 3 -f -o gpurun_out/prof_f $S > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
tail -n 3 gpurun_out/tests_gpu.log; cut -c1-400 gpurun_out/bench_n1.json; cut -c1-300 gpurun_out/bench_ref.json
