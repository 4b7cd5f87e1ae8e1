#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/tests_gpu.log 2>&1; echo "tests rc=$?"
tail -n 15 gpurun_out/tests_gpu.log
timeout 200 python tools/e2e_trace.py 2>&1 | tail -12
