#!/bin/bash
# one GPU call: variant sweep, single-frame configs, parity tests in the riskiest and the default mode
mkdir -p gpurun_out
python tools/variants_prof.py > gpurun_out/variants.log 2>&1
echo "variants rc=$?"
BVX_PEEK=1 BVX_SMALL_TABLE=1 python tools/cfg_prof.py > gpurun_out/cfg_peek_small.log 2>&1
python tools/cfg_prof.py > gpurun_out/cfg_default.log 2>&1
BVX_PEEK=1 BVX_SMALL_TABLE=1 timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/tests_peek_small.log 2>&1
echo "tests peek+small rc=$?"
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/tests_default.log 2>&1
echo "tests default rc=$?"
tail -n 30 gpurun_out/variants.log
tail -n 5 gpurun_out/cfg_peek_small.log gpurun_out/cfg_default.log gpurun_out/tests_peek_small.log gpurun_out/tests_default.log
