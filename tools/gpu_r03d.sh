#!/bin/bash
# r03d: allocation-free entry point (alone in a fresh process, then after the whole suite), identity factor removed from the noise
mkdir -p gpurun_out
timeout -k 10 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r03d_pytest_noalloc_alone.txt 2>&1; echo "alone rc=$?"; tail -15 gpurun_out/r03d_pytest_noalloc_alone.txt | cut -c1-300
timeout -k 10 1200 python -m pytest tests -x -q -m gpu > gpurun_out/r03d_pytest_gpu.txt 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/r03d_pytest_gpu.txt | cut -c1-300
timeout -k 10 300 python bench.py --no-configs --steps 10 --warmup 3 > gpurun_out/r03d_bench.json 2> gpurun_out/r03d_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
j = json.loads(open("gpurun_out/r03d_bench.json").read().strip().splitlines()[-1])
print("ms/step", j["ms_per_step"], "wp_ms", j["roofline"]["avg_launch_ms"], "frac", j["roofline"]["frac"], j["roofline"]["other_kernels"], "e2e", j["e2e"]["ms_per_step"])
PY
