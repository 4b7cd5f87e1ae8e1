#!/bin/bash
# round-end evidence: GPU parity tests, bench lines, ncu launch list, ncu full set of the four pipeline kernels
mkdir -p gpurun_out
S="--steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/tests_gpu.log 2>&1; echo "tests rc=$?"
timeout 600 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 300 python bench.py $S > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py $S > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_insert|k_slot|k_gather|k_verify" -s 8 -c 4 -f -o gpurun_out/prof_f python bench.py $S > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
tail -n 3 gpurun_out/tests_gpu.log; cut -c1-400 gpurun_out/bench_n1.json; cut -c1-300 gpurun_out/bench_ref.json
