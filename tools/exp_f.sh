#!/bin/bash
# one GPU call: cfg4 per-kernel times, single-frame configs, GPU parity tests
mkdir -p gpurun_out
timeout 200 python tools/variants_prof.py > gpurun_out/variants.log 2>&1
timeout 200 python tools/cfg_prof.py > gpurun_out/cfg_default.log 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/tests_default.log 2>&1
echo "tests rc=$?"
cat gpurun_out/variants.log gpurun_out/cfg_default.log; tail -n 4 gpurun_out/tests_default.log
