#!/bin/bash
# parity suite + short default bench (no CPU leg).  Usage: gpurun --timeout 900 -- 'bash tools/gpu_quick2.sh <tag>'
tag=${1:-q}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$tag.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_$tag.log
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err; echo "bench rc=$?"
python - <<PY
import json
j = json.load(open("gpurun_out/bench_$tag.json"))
print('ms/step', round(j['ms_per_step'],3), 'wp_ms', round(j['roofline']['avg_launch_ms'],3), 'mapgen_ms', round(j['roofline']['other_kernels']['k_mapgen_grid_avg_ms'],3), 'frac', round(j['roofline']['frac'],4), 'e2e ms', j['e2e'] and round(j['e2e']['ms_per_step'],2))
PY
tail -5 gpurun_out/bench_$tag.err
