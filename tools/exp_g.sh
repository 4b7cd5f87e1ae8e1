#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/tests_gpu.log 2>&1; echo "tests rc=$?"
tail -n 15 gpurun_out/tests_gpu.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_g.json 2> gpurun_out/bench_g.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_g.json').read().strip().splitlines()[-1])
for k in ('e2e','e2e_device_rows','e2e_full_upload'):
    e=d[k]; print(k, e['value'], e['ms_per_step'], e['h2d_bytes_per_step'], e['d2h_bytes_per_step'], e.get('host_threads'))
print(d['value'], d['ms_per_step'])
PY
nproc; lscpu | grep -E "Model name|Socket|NUMA node\(s\)|^CPU\(s\)"
