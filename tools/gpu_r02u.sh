#!/bin/bash
# r02u: relaxation step of the one-CTA kernel without a branch around the atomics, weights loaded next to the label
mkdir -p gpurun_out
timeout -k 10 900 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_r02u.txt 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_r02u.txt
timeout -k 10 300 python bench.py --no-configs --steps 10 --warmup 3 > gpurun_out/bench_r02u.json 2> gpurun_out/bench_r02u.err; echo "bench rc=$?"
python - <<'PY'
import json
j = json.loads(open("gpurun_out/bench_r02u.json").read().strip().splitlines()[-1])
print("ms/step", j["ms_per_step"], "wp_ms", j["roofline"]["avg_launch_ms"], "frac", j["roofline"]["frac"], j["roofline"]["other_kernels"], "e2e", j["e2e"]["value"])
PY
MTD_LIB=$PWD/mazetd_b200/variants/libmazetd_b200_timing.so timeout -k 10 200 python bench.py --no-configs --no-e2e --steps 3 --warmup 3 > gpurun_out/phases_r02u.txt 2>&1; echo "timing rc=$?"
grep -E "MAPGEN|PHASES|TIMING" gpurun_out/phases_r02u.txt | head -3
