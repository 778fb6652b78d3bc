#!/bin/bash
# r02p: larger scripted-events test, headline kernel after removing the wide-bucket option, phase split of k_waypoints_small
mkdir -p gpurun_out
timeout -k 10 400 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "larger_size or incremental or scripted" > gpurun_out/pytest_r02p.txt 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_r02p.txt
timeout -k 10 300 python bench.py --no-configs --steps 10 --warmup 3 > gpurun_out/bench_r02p.json 2> gpurun_out/bench_r02p.err; echo "bench rc=$?"
python - <<'PY'
import json
j = json.loads(open("gpurun_out/bench_r02p.json").read().strip().splitlines()[-1])
print("ms/step", j["ms_per_step"], "wp_ms", j["roofline"]["avg_launch_ms"], "frac", j["roofline"]["frac"], "e2e", j["e2e"]["value"])
PY
MTD_LIB=$PWD/mazetd_b200/variants/libmazetd_b200_timing.so timeout -k 10 300 python bench.py --no-configs --steps 3 --warmup 3 > gpurun_out/phases_r02p.txt 2>&1; echo "timing rc=$?"
grep -E "PHASES|TIMING" gpurun_out/phases_r02p.txt | tail -4
