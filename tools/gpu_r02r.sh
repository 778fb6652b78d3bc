#!/bin/bash
# r02r: decoupled walker warps + 4-layer write-out; parity tests, headline, phase split of both small-map kernels
mkdir -p gpurun_out
timeout -k 10 900 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_r02r.txt 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_r02r.txt
timeout -k 10 300 python bench.py --no-configs --steps 10 --warmup 3 > gpurun_out/bench_r02r.json 2> gpurun_out/bench_r02r.err; echo "bench rc=$?"
python - <<'PY'
import json
j = json.loads(open("gpurun_out/bench_r02r.json").read().strip().splitlines()[-1])
print("ms/step", j["ms_per_step"], "wp_ms", j["roofline"]["avg_launch_ms"], "frac", j["roofline"]["frac"], "e2e", j["e2e"]["value"])
PY
MTD_LIB=$PWD/mazetd_b200/variants/libmazetd_b200_timing.so timeout -k 10 200 python bench.py --no-configs --steps 3 --warmup 3 > gpurun_out/phases_r02r.txt 2>&1; echo "timing rc=$?"
grep -E "MAPGEN|PHASES" gpurun_out/phases_r02r.txt | tail -4
